// noise.cu - synthetic fBm DEM generator (the input of every benchmark config).
//
// Reference: source/soillib/op/noise.hpp:14-56 - a serial host loop over
// FastNoiseLite 1.0.1 (vendored third-party header, not part of soillib's own
// code) configured as OpenSimplex2 + FBm; value(x,y) = GetNoise(x/ext.x,
// y/ext.y, seed) where `seed` is the z coordinate of 3-D noise and the integer
// seed stays at FastNoiseLite's default 1337.
//
// Restated from the published OpenSimplex2 algorithm as host+device code so the
// same bits can be produced by one kernel on the GPU (an 8192^2 DEM in well
// under a millisecond instead of ~10 s of serial host work) or on the host.
// Pinned by tests/golden/fnl_*.npy generated from the vendored header.
#include "common.cuh"
#include <vector>

namespace sb {
namespace {

// the 64 gradient directions: five repeats of the 12 cube-edge midpoints, then four extras
__host__ __device__ inline void grad3(int idx, float& gx, float& gy, float& gz) {
  int e = idx < 60 ? idx % 12 : -1;
  if (e < 0) {
    const int j = idx - 60;  // (1,1,0) (0,-1,1) (-1,1,0) (0,-1,-1)
    gx = j == 0 ? 1.0f : (j == 2 ? -1.0f : 0.0f);
    gy = (j & 1) ? -1.0f : 1.0f;
    gz = j == 1 ? 1.0f : (j == 3 ? -1.0f : 0.0f);
    return;
  }
  const float s0 = (e & 1) ? -1.0f : 1.0f;  // sign of the first non-zero component
  const float s1 = (e & 2) ? -1.0f : 1.0f;  // sign of the second
  const int grp = e >> 2;                   // 0: (0,a,b)  1: (a,0,b)  2: (a,b,0)
  gx = grp == 0 ? 0.0f : s0;
  gy = grp == 0 ? s0 : (grp == 1 ? 0.0f : s1);
  gz = grp == 2 ? 0.0f : s1;
}

__host__ __device__ inline float grad_coord(int seed, int xp, int yp, int zp, float xd, float yd, float zd) {
  unsigned hash = (unsigned)(seed ^ xp ^ yp ^ zp) * 0x27d4eb2du;
  int h = (int)hash;
  h ^= h >> 15;
  h &= 63 << 2;
  float gx, gy, gz;
  grad3(h >> 2, gx, gy, gz);
  return xd * gx + yd * gy + zd * gz;
}

__host__ __device__ inline int fast_round(float f) { return f >= 0 ? (int)(f + 0.5f) : (int)(f - 0.5f); }

__host__ __device__ inline float simplex3(int seed, float x, float y, float z) {
  const unsigned PX = 501125321u, PY = 1136930381u, PZ = 1720413743u;
  int i = fast_round(x), j = fast_round(y), k = fast_round(z);
  float x0 = (float)(x - i), y0 = (float)(y - j), z0 = (float)(z - k);
  int xs = (int)(-1.0f - x0) | 1, ys = (int)(-1.0f - y0) | 1, zs = (int)(-1.0f - z0) | 1;
  float ax0 = xs * -x0, ay0 = ys * -y0, az0 = zs * -z0;
  unsigned iu = (unsigned)i * PX, ju = (unsigned)j * PY, ku = (unsigned)k * PZ;
  float value = 0;
  float a = (0.6f - x0 * x0) - (y0 * y0 + z0 * z0);
  for (int l = 0;; l++) {
    if (a > 0) value += (a * a) * (a * a) * grad_coord(seed, (int)iu, (int)ju, (int)ku, x0, y0, z0);
    float b = a + 1;
    unsigned i1 = iu, j1 = ju, k1 = ku;
    float x1 = x0, y1 = y0, z1 = z0;
    if (ax0 >= ay0 && ax0 >= az0) {
      x1 += xs; b -= xs * 2 * x1; i1 -= (unsigned)xs * PX;
    } else if (ay0 > ax0 && ay0 >= az0) {
      y1 += ys; b -= ys * 2 * y1; j1 -= (unsigned)ys * PY;
    } else {
      z1 += zs; b -= zs * 2 * z1; k1 -= (unsigned)zs * PZ;
    }
    if (b > 0) value += (b * b) * (b * b) * grad_coord(seed, (int)i1, (int)j1, (int)k1, x1, y1, z1);
    if (l == 1) break;
    ax0 = 0.5f - ax0; ay0 = 0.5f - ay0; az0 = 0.5f - az0;
    x0 = xs * ax0; y0 = ys * ay0; z0 = zs * az0;
    a += (0.75f - ax0) - (ay0 + az0);
    iu += (unsigned)(xs >> 1) & PX; ju += (unsigned)(ys >> 1) & PY; ku += (unsigned)(zs >> 1) & PZ;
    xs = -xs; ys = -ys; zs = -zs;
    seed = ~seed;
  }
  return value * 32.69428253173828125f;
}

struct fbm_t {
  float frequency, gain, lacunarity, bounding;
  int octaves;
};

__host__ __device__ inline float fbm(const fbm_t& f, float x, float y, float z) {
  x *= f.frequency; y *= f.frequency; z *= f.frequency;
  const float R3 = (float)(2.0 / 3.0);  // OpenSimplex2's default 3-D rotation
  const float r = (x + y + z) * R3;
  x = r - x; y = r - y; z = r - z;
  int seed = 1337;
  float sum = 0, amp = f.bounding;
  for (int i = 0; i < f.octaves; i++) {
    const float noise = simplex3(seed++, x, y, z);
    sum += noise * amp;
    amp *= 1.0f + 0.0f * ((noise + 1) * 0.5f - 1.0f);  // weighted strength 0
    x *= f.lacunarity; y *= f.lacunarity; z *= f.lacunarity;
    amp *= f.gain;
  }
  return sum;
}

__global__ void __launch_bounds__(256) k_noise(float* __restrict__ out, int nx, int ny, fbm_t f, float seed, float ex,
                                               float ey) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  out[n] = fbm(f, x / ex, y / ey, seed);
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" int soilb200_noise_f32(float* out, int nx, int ny, float frequency, int octaves, float gain,
                                  float lacunarity, float seed, float ext_x, float ext_y, int on_device,
                                  void* stream) {
  if (!out || nx <= 0 || ny <= 0 || octaves <= 0) return SOILB200_E_ARG;
  fbm_t f;
  f.frequency = frequency; f.gain = gain; f.lacunarity = lacunarity; f.octaves = octaves;
  float g = gain < 0 ? -gain : gain, amp = g, ampFractal = 1.0f;
  for (int i = 1; i < octaves; i++) { ampFractal += amp; amp *= g; }
  f.bounding = 1 / ampFractal;
  if (on_device) {
    const long long N = (long long)nx * ny;
    k_noise<<<blocks_for(N, 256), 256, 0, (cudaStream_t)stream>>>(out, nx, ny, f, seed, ext_x, ext_y);
    count_launch();
    return check_launch();
  }
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) out[(size_t)x * ny + y] = fbm(f, x / ext_x, y / ext_y, seed);
  return SOILB200_OK;
}
