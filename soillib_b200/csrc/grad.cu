// grad.cu - SURVEY.md section 8(f) row 1: gradient, negslope, laplacian (grad.cu:22-206 of the reference),
// slope along the flow graph (graph.cu:270-311) and the separable Gaussian blur (filter.cu:24-91).
//
// All are streaming stencils over float32 maps (HBM-bound: 12, 8, 8-16, 12 and 2 x 8 bytes per cell).  One thread
// per cell in 32 x 8 blocks with the fast axis along the warp, so every row of a stencil is a coalesced load and the
// vertical taps hit lines the neighbouring warps of the block just fetched.  Arithmetic follows the reference's
// operand order with IEEE divisions and no FMA contraction (-fmad=false), so results are bit-identical to the CPU
// oracle (the blur up to its __expf weights) and to the reference's kernels built with -fmad=false.
#include "common.cuh"

namespace sb {
namespace {

struct cell {
  int x, y;
  long long n;
  bool in;
};
__device__ __forceinline__ cell my_cell(int nx, int ny) {
  cell c;
  c.y = blockIdx.x * 32 + (threadIdx.x & 31);
  c.x = blockIdx.y * 8 + (threadIdx.x >> 5);
  c.in = c.x < nx && c.y < ny;
  c.n = (long long)c.x * ny + c.y;
  return c;
}
inline dim3 cell_grid(int nx, int ny) { return dim3((ny + 31) / 32, (nx + 7) / 8); }

// the five heights of the reference's loaders: NaN outside the map (grad.cu:33-38)
struct h5 {
  float c, n0, p0, zn, zp;
};
__device__ __forceinline__ h5 load5(const float* __restrict__ h, const cell& c, int nx, int ny) {
  h5 w;
  w.c = __ldg(h + c.n);
  w.n0 = c.x > 0 ? __ldg(h + c.n - ny) : CUDART_NAN_F;
  w.p0 = c.x + 1 < nx ? __ldg(h + c.n + ny) : CUDART_NAN_F;
  w.zn = c.y > 0 ? __ldg(h + c.n - 1) : CUDART_NAN_F;
  w.zp = c.y + 1 < ny ? __ldg(h + c.n + 1) : CUDART_NAN_F;
  return w;
}

// grad.cu:22-89
__global__ void __launch_bounds__(256) k_gradient(const float* __restrict__ h, float2* __restrict__ out, int nx, int ny,
                                                  float sx, float sy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
  const h5 w = load5(h, c, nx, ny);
  const float gxn = (w.c - w.n0) / sx, gyn = (w.c - w.zn) / sy;
  const float gxp = (w.p0 - w.c) / sx, gyp = (w.zp - w.c) / sy;
  float gx = 0.5f * (w.p0 - w.n0) / sx;
  float gy = 0.5f * (w.zp - w.zn) / sy;
  if (gx != gx) gx = gxn;
  if (gx != gx) gx = gxp;
  if (gx != gx) gx = 0.0f;
  if (gy != gy) gy = gyn;
  if (gy != gy) gy = gyp;
  if (gy != gy) gy = 0.0f;
  out[c.n] = make_float2(gx, gy);
}

// grad.cu:102-130; glm::max(a, b) = (a < b) ? b : a
__device__ __forceinline__ float glm_max(float a, float b) { return a < b ? b : a; }
__global__ void __launch_bounds__(256) k_negslope(const float* __restrict__ h, float* __restrict__ out, int nx, int ny,
                                                  float sx, float sy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
  const h5 w = load5(h, c, nx, ny);
  float gx = 0.0f, gy = 0.0f;
  if (w.n0 == w.n0) gx = glm_max(gx, (w.c - w.n0) / sx);
  if (w.p0 == w.p0) gx = glm_max(gx, (w.c - w.p0) / sx);
  if (w.zn == w.zn) gy = glm_max(gy, (w.c - w.zn) / sy);
  if (w.zp == w.zp) gy = glm_max(gy, (w.c - w.zp) / sy);
  out[c.n] = len2(gx, gy);
}

// grad.cu:137-176, D interleaved components per cell
template <int D>
__global__ void __launch_bounds__(256) k_laplacian(const float* __restrict__ in, float* __restrict__ out, int nx, int ny,
                                                   float sx, float sy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
  const float hx = (1.0f / sx / sx);
  const float hy = (1.0f / sy / sy);
  const bool xm = c.x > 0, xp = c.x + 1 < nx, ym = c.y > 0, yp = c.y + 1 < ny;
#pragma unroll
  for (int d = 0; d < D; ++d) {
    auto at = [&](bool ok, long long off) { return ok ? __ldg(in + (c.n + off) * D + d) : __ldg(in + c.n * D + d); };
    const float v00 = __ldg(in + c.n * D + d);
    const float vn0 = at(xm, -ny), vp0 = at(xp, ny), v0n = at(ym, -1), v0p = at(yp, 1);
    const float vnn = at(xm && ym, -ny - 1), vpp = at(xp && yp, ny + 1), vpn = at(xp && ym, ny - 1), vnp = at(xm && yp, -ny + 1);
    const float LH = (vn0 - v00) * hx + (vp0 - v00) * hx + (v0n - v00) * hy + (v0p - v00) * hy;
    const float LD = 0.5f * (vnn - v00) * hx + 0.5f * (vpp - v00) * hx + 0.5f * (vpn - v00) * hy + 0.5f * (vnp - v00) * hy;
    out[c.n * D + d] = 0.5f * LH + 0.5f * LD;
  }
}

// graph.cu:270-296
__global__ void __launch_bounds__(256) k_slope(const float* __restrict__ h, const int32_t* __restrict__ flow,
                                               float* __restrict__ out, int nx, int ny, float sx, float sy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
  const int next = __ldg(flow + c.n);
  if (next < 0 || next == (int)c.n || next >= nx * ny) {  // no receiver (the last test only keeps a corrupt graph from faulting)
    out[c.n] = 0.0f;
    return;
  }
  const float dx = sx * ((float)(next / ny) - (float)c.x), dy = sy * ((float)(next % ny) - (float)c.y);
  out[c.n] = (__ldg(h + next) - __ldg(h + c.n)) / len2(dx, dy);
}

// filter.cu:24-57, one pass: 33 taps along dim 0 (xdir) or dim 1, indices clamped at the map edge, weights not
// renormalised there.  The map is handled as nx rows of ny * channels floats; the 33 weights (filter.cu:45-46,
// the same expression for every cell) are computed once per block.
constexpr int kBlurWindow = 16;
__global__ void __launch_bounds__(256) k_blur(const float* __restrict__ src, float* __restrict__ dst, int nx, int ny,
                                              int channels, float sigma, bool xdir) {
  __shared__ float weight[2 * kBlurWindow + 1];
  if (threadIdx.x <= 2 * kBlurWindow) {
    const int k = (int)threadIdx.x - kBlurWindow;
    const float Z = sqrtf(2.0f * 3.14159265f) * sigma;
    weight[threadIdx.x] = __expf(-0.5f * (k / sigma) * (k / sigma)) / Z;
  }
  __syncthreads();
  const int W = ny * channels;
  const int col = blockIdx.x * 32 + (threadIdx.x & 31);  // y * channels + ch
  const int x = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= nx || col >= W) return;
  const int y = col / channels, ch = col - y * channels;
  float val = 0.0f;
#pragma unroll 11
  for (int k = -kBlurWindow; k <= kBlurWindow; ++k) {
    int px = x + (xdir ? k : 0), py = y + (xdir ? 0 : k);
    px = min(max(px, 0), nx - 1);
    py = min(max(py, 0), ny - 1);
    val += __ldg(src + (size_t)px * W + (size_t)py * channels + ch) * weight[k + kBlurWindow];
  }
  dst[(size_t)x * W + col] = val;
}

inline bool bad2(const float* s) { return !s || !(s[0] == s[0]) || !(s[1] == s[1]); }

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

int soilb200_gradient_f32(const float* height, float* out, int nx, int ny, const float scale[2], void* stream) {
  if (!height || !out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  k_gradient<<<cell_grid(nx, ny), 256, 0, (cudaStream_t)stream>>>(height, (float2*)out, nx, ny, scale[0], scale[1]);
  count_launch();
  return check_launch();
}

int soilb200_negslope_f32(const float* height, float* out, int nx, int ny, const float scale[2], void* stream) {
  if (!height || !out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  k_negslope<<<cell_grid(nx, ny), 256, 0, (cudaStream_t)stream>>>(height, out, nx, ny, scale[0], scale[1]);
  count_launch();
  return check_launch();
}

int soilb200_laplacian_f32(const float* in, float* out, int nx, int ny, int channels, const float scale[2],
                           void* stream) {
  if (!in || !out || in == out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if (channels != 1 && channels != 2) return SOILB200_E_ARG;  // the reference instantiates D = 1, 2 (grad.cu:193-199)
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  if (channels == 1) k_laplacian<1><<<cell_grid(nx, ny), 256, 0, s>>>(in, out, nx, ny, scale[0], scale[1]);
  else k_laplacian<2><<<cell_grid(nx, ny), 256, 0, s>>>(in, out, nx, ny, scale[0], scale[1]);
  count_launch();
  return check_launch();
}

int soilb200_slope_f32(const float* height, const int32_t* flow, float* out, int nx, int ny, const float scale[2],
                       void* stream) {
  if (!height || !flow || !out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  k_slope<<<cell_grid(nx, ny), 256, 0, (cudaStream_t)stream>>>(height, flow, out, nx, ny, scale[0], scale[1]);
  count_launch();
  return check_launch();
}

int soilb200_gaussian_blur_f32(float* inout, float* tmp, int nx, int ny, int channels, float sigma, void* stream) {
  if (!inout || !tmp || inout == tmp || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if (channels != 1 && channels != 2) return SOILB200_E_ARG;  // filter.cu:79-88
  if ((long long)nx * ny * channels > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  const dim3 grid((ny * channels + 31) / 32, (nx + 7) / 8);
  k_blur<<<grid, 256, 0, s>>>(inout, tmp, nx, ny, channels, sigma, true);
  k_blur<<<grid, 256, 0, s>>>(tmp, inout, nx, ny, channels, sigma, false);
  count_launch(2);
  return check_launch();
}

}  // extern "C"
