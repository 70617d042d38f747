// path.cu - SURVEY.md section 8(f) row 3: soil::solve_uniform (reference path.cu:51-219, sample.hpp:72-186), the
// grid-free Monte-Carlo estimator for a linear conservation law: uniformly started samples follow a bilinearly
// interpolated flow field and credit every cell they enter with their source rate, attenuated by the decay
// collected along the way; one dense pass normalises the estimate.
//
// Same particle engine as the erosion transport (transport.cu): one thread per sample, blocks of 64 so that the
// few long-lived warps spread over all SMs, a serial chain of scattered gathers per crossing (four 8-byte flow
// vectors for the bilinear support + the decay of the cell) and one or two float reductions per crossing - bound by
// DRAM latency and random-sector rate, not by bandwidth.  The walk does not depend on the attenuation, so the
// trajectories are bit-identical to the CPU oracle's; the deposits differ by __expf and the order of the atomics.
#include "erosion_math.cuh"
#include <curand_kernel.h>

namespace sb {
namespace {

constexpr int kSampleBlock = 64;

// float -> int as the reference's glm conversions do on the device (truncation; cvt.rzi maps NaN to 0)
__device__ __forceinline__ int trunc_i(float f) { return __float2int_rz(f); }

__device__ __forceinline__ bool oob_i(float px, float py, int nx, int ny) {  // shape.oob(ivec2(pos))
  const int x = trunc_i(px), y = trunc_i(py);
  return x < 0 || y < 0 || x >= nx || y >= ny;
}

// sample_t<vec2,2,1>::gather(view, res, pos).val(), sample.hpp:151-186 and :86-88 (row-major view overload)
__device__ __forceinline__ float2 flow_at(const float2* __restrict__ flow, int nx, int ny, float px, float py) {
  if (px < 0 || py < 0 || px > nx - 1 || py > ny - 1) return make_float2(CUDART_NAN_F, CUDART_NAN_F);
  const int ix = trunc_i(px), iy = trunc_i(py);
  float wx = px - floorf(px), wy = py - floorf(py);
  int i00 = ix * ny + iy, i01 = ix * ny + (iy + 1), i10 = (ix + 1) * ny + iy, i11 = (ix + 1) * ny + (iy + 1);
  if (px + 1 > nx - 1) { wx = 0; i10 = 0; i11 = 0; }
  if (py + 1 > ny - 1) { wy = 0; i01 = 0; i11 = 0; }
  const float2 h00 = __ldg(flow + i00), h01 = __ldg(flow + i01), h10 = __ldg(flow + i10), h11 = __ldg(flow + i11);
  // M * (1, t) = (1 + (-1) t, 0 + t), as the matrix product forms the weights
  const float ay = 1.0f * 1.0f + -1.0f * wy, by = 0.0f * 1.0f + 1.0f * wy;
  const float2 l0 = make_float2(ay * h00.x + by * h01.x, ay * h00.y + by * h01.y);
  const float2 l1 = make_float2(ay * h10.x + by * h11.x, ay * h10.y + by * h11.y);
  const float ax = 1.0f * 1.0f + -1.0f * wx, bx = 0.0f * 1.0f + 1.0f * wx;
  return make_float2(ax * l0.x + bx * l1.x, ax * l0.y + bx * l1.y);
}

// path.cu:51-137
template <int K>
__global__ void __launch_bounds__(kSampleBlock) k_solve_uniform(float* __restrict__ flux, const float2* __restrict__ flow,
                                                                const float* __restrict__ source,
                                                                const float* __restrict__ decay,
                                                                curandStateXORWOW_t* __restrict__ rng, long long P,
                                                                int nx, int ny, float sx, float sy, float epsilon,
                                                                float maxstep) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= P) return;
  curandStateXORWOW_t rs = rng[n];
  float px = curand_uniform(&rs) * float(nx);
  float py = curand_uniform(&rs) * float(ny);
  rng[n] = rs;
  // a draw of exactly 1.0 starts outside the map; the reference reads out of bounds there
  if (oob_i(px, py, nx, ny)) return;
  int ind = trunc_i(px) * ny + trunc_i(py);

  const float L = len2(sx, sy);
  const float A = sx * sy;
  const float Pr = 1.0f / (A * (float)((size_t)nx * (size_t)ny));
  float S[K];
#pragma unroll
  for (int c = 0; c < K; ++c) S[c] = __ldg(source + (size_t)ind * K + c) / Pr;
  const float lenS = K == 1 ? sqrtf(S[0] * S[0]) : len2(S[0], S[K - 1]);
  if (lenS < epsilon) return;

  float att = 1.0f;
  int step = 0;
  while (!oob_i(px, py, nx, ny) && epsilon < fabsf(att) && (float)(++step) < maxstep) {
    const int nind = trunc_i(px) * ny + trunc_i(py);
    if (nind != ind) {
      ind = nind;
#pragma unroll
      for (int c = 0; c < K; ++c) atomicAdd(flux + (size_t)ind * K + c, S[c] * att);
    }
    const float dec = __ldg(decay + ind);  // issued with the flow gathers, used after the step
    const float2 v = flow_at(flow, nx, ny, px, py);
    const float v_len = len2(v.x, v.y);
    if (v_len < epsilon) break;
    const float ux = v.x / v_len, uy = v.y / v_len;
    const float st = stepsize(px, py, ux, uy);  // path.cu:26-49 == erosion_map.cu:56-78
    px += st * ux;
    py += st * uy;
    const float dlambda = st * L / v_len;
    att *= __expf(-dlambda * dec);
  }
}

// path.cu:139-167: (source * A + flux / count) / (|v.x * scale.y| + |v.y * scale.x|), in place
template <int K>
__global__ void __launch_bounds__(256) k_solve_normalize(float* __restrict__ flux, const float2* __restrict__ flow,
                                                         const float* __restrict__ source, long long N, float sx,
                                                         float sy, float count) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float2 v = __ldcs(flow + n);
  const float A = sx * sy;
  const float norm = fabsf(v.x * sy) + fabsf(v.y * sx);
#pragma unroll
  for (int c = 0; c < K; ++c) {
    const size_t i = (size_t)n * K + c;
    flux[i] = (__ldcs(source + i) * A + flux[i] / count) / norm;
  }
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

int soilb200_solve_uniform_f32(const float* flow, const float* source, const float* decay, void* rng,
                               int64_t n_samples, float* flux, int nx, int ny, int components, const float scale[2],
                               int64_t count, void* stream) {
  if (!flow || !source || !decay || !rng || !flux || !scale || nx <= 0 || ny <= 0 || n_samples < 0 || count <= 0)
    return SOILB200_E_ARG;
  if (components != 1 && components != 2) return SOILB200_E_ARG;  // path.cu:198-209
  if ((long long)nx * ny * components > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  const long long N = (long long)nx * ny;
  cudaError_t e = cudaMemsetAsync(flux, 0, (size_t)N * components * sizeof(float), s);  // silt::set(flux, 0), path.cu:190
  if (e != cudaSuccess) return (int)e;
  const float epsilon = 1E-16f;               // path.cu:193
  const float maxstep = (float)(nx + ny);     // path.cu:194
  if (n_samples > 0) {
    const unsigned nb = blocks_for(n_samples, kSampleBlock);
    if (components == 1)
      k_solve_uniform<1><<<nb, kSampleBlock, 0, s>>>(flux, (const float2*)flow, source, decay, (curandStateXORWOW_t*)rng,
                                                     n_samples, nx, ny, scale[0], scale[1], epsilon, maxstep);
    else
      k_solve_uniform<2><<<nb, kSampleBlock, 0, s>>>(flux, (const float2*)flow, source, decay, (curandStateXORWOW_t*)rng,
                                                     n_samples, nx, ny, scale[0], scale[1], epsilon, maxstep);
    count_launch();
  }
  if (components == 1)
    k_solve_normalize<1><<<blocks_for(N, 256), 256, 0, s>>>(flux, (const float2*)flow, source, N, scale[0], scale[1], (float)count);
  else
    k_solve_normalize<2><<<blocks_for(N, 256), 256, 0, s>>>(flux, (const float2*)flow, source, N, scale[0], scale[1], (float)count);
  count_launch();
  return check_launch();
}

}  // extern "C"
