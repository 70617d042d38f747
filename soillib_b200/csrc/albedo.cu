// albedo.cu - SURVEY.md section 8(f) row 4: the albedo generators of the reference (erosion.cu:759-919):
// albedo_layer (bedrock / sediment colour blend by sediment thickness), albedo_stratum (alternating strata colours
// by depth below the accumulated uplift) and albedo_discharge (extinction blend towards a water colour).
//
// Per-cell streaming over interleaved (nx, ny, 3) colour maps.  A thread handles 4 cells, so that the 12 colour
// floats of its cells move as three 16-byte accesses per map instead of twelve 4-byte ones at a 12-byte stride;
// a scalar tail / fallback covers cell counts that are not a multiple of 4 and unaligned bases.  Arithmetic in the
// reference's operand order without FMA contraction: bit-identical to the CPU oracle except for the __expf of
// albedo_discharge, and to the reference's kernels built with -fmad=false.
#include "common.cuh"
#include <initializer_list>

namespace sb {
namespace {

struct rgb {
  float r, g, b;
};

// 12 floats = the colours of 4 consecutive cells
struct quad12 {
  float v[12];
};
__device__ __forceinline__ quad12 load12(const float* __restrict__ p) {
  quad12 q;
  const float4 a = __ldcs(reinterpret_cast<const float4*>(p)), b = __ldcs(reinterpret_cast<const float4*>(p) + 1),
               c = __ldcs(reinterpret_cast<const float4*>(p) + 2);
  q.v[0] = a.x; q.v[1] = a.y; q.v[2] = a.z; q.v[3] = a.w;
  q.v[4] = b.x; q.v[5] = b.y; q.v[6] = b.z; q.v[7] = b.w;
  q.v[8] = c.x; q.v[9] = c.y; q.v[10] = c.z; q.v[11] = c.w;
  return q;
}
__device__ __forceinline__ void store12(float* __restrict__ p, const quad12& q) {
  __stcs(reinterpret_cast<float4*>(p), make_float4(q.v[0], q.v[1], q.v[2], q.v[3]));
  __stcs(reinterpret_cast<float4*>(p) + 1, make_float4(q.v[4], q.v[5], q.v[6], q.v[7]));
  __stcs(reinterpret_cast<float4*>(p) + 2, make_float4(q.v[8], q.v[9], q.v[10], q.v[11]));
}

// erosion.cu:773-777 for one colour component
__device__ __forceinline__ float layer_mix(float bedrock, float sediment, float shift, float blend) {
  const float colorSediment = fminf(sediment + shift, 1.0f);  // __mmin(1.0f, .), erosion_map.cu:88-97
  return blend * bedrock + (1.0f - blend) * colorSediment;
}
__device__ __forceinline__ float layer_blend(float sedimentHeight, float scaleSediment) {
  return 1.0f / (1.0f + scaleSediment * sedimentHeight);
}

template <bool VEC>
__global__ void __launch_bounds__(256) k_albedo_layer(float* __restrict__ albedo, const float* __restrict__ bedrock,
                                                      const float* __restrict__ sediment,
                                                      const float2* __restrict__ layers, long long n,
                                                      float scaleSediment, rgb shift) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const float sh[3] = {shift.r, shift.g, shift.b};
  if (VEC) {
    const long long i0 = 4 * t;
    if (i0 >= n) return;
    const quad12 cb = load12(bedrock + 3 * i0), cs = load12(sediment + 3 * i0);
    const float4 la = __ldcs(reinterpret_cast<const float4*>(layers + i0)),
                 lb = __ldcs(reinterpret_cast<const float4*>(layers + i0) + 1);
    const float sed[4] = {la.y, la.w, lb.y, lb.w};
    quad12 o;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float blend = layer_blend(sed[j], scaleSediment);
#pragma unroll
      for (int c = 0; c < 3; ++c) o.v[3 * j + c] = layer_mix(cb.v[3 * j + c], cs.v[3 * j + c], sh[c], blend);
    }
    store12(albedo + 3 * i0, o);
  } else {
    if (t >= n) return;
    const float blend = layer_blend(__ldg(&layers[t].y), scaleSediment);
#pragma unroll
    for (int c = 0; c < 3; ++c)
      albedo[3 * t + c] = layer_mix(__ldg(bedrock + 3 * t + c), __ldg(sediment + 3 * t + c), sh[c], blend);
  }
}

// erosion.cu:808-824
__device__ __forceinline__ bool stratum_is_a(float uplift, float bedrockHeight, float age_uplift, float sz, float freq) {
  const float shift = age_uplift * uplift;
  const float depth = fmaxf(shift - bedrockHeight * sz, 0.0f);
  const int index = floorf(depth / freq);
  return index % 2 == 0;
}

template <bool VEC>
__global__ void __launch_bounds__(256) k_albedo_stratum(float* __restrict__ albedoBedrock,
                                                        const float* __restrict__ uplift,
                                                        const float2* __restrict__ layers, long long n,
                                                        float age_uplift, float sz, float freq, rgb A, rgb B) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (VEC) {
    const long long i0 = 4 * t;
    if (i0 >= n) return;
    const float4 u = __ldcs(reinterpret_cast<const float4*>(uplift + i0));
    const float4 la = __ldcs(reinterpret_cast<const float4*>(layers + i0)),
                 lb = __ldcs(reinterpret_cast<const float4*>(layers + i0) + 1);
    const float up[4] = {u.x, u.y, u.z, u.w}, bed[4] = {la.x, la.z, lb.x, lb.z};
    quad12 o;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool a = stratum_is_a(up[j], bed[j], age_uplift, sz, freq);
      o.v[3 * j] = a ? A.r : B.r; o.v[3 * j + 1] = a ? A.g : B.g; o.v[3 * j + 2] = a ? A.b : B.b;
    }
    store12(albedoBedrock + 3 * i0, o);
  } else {
    if (t >= n) return;
    const bool a = stratum_is_a(__ldg(uplift + t), __ldg(&layers[t].x), age_uplift, sz, freq);
    albedoBedrock[3 * t] = a ? A.r : B.r; albedoBedrock[3 * t + 1] = a ? A.g : B.g; albedoBedrock[3 * t + 2] = a ? A.b : B.b;
  }
}

// erosion.cu:871-874
__device__ __forceinline__ float discharge_blend(float discharge, float extinction, float scale) {
  const float value = fmaxf(0.0f, discharge);
  return scale * (1.0f - __expf(-extinction * value));
}

template <bool VEC>
__global__ void __launch_bounds__(256) k_albedo_discharge(float* __restrict__ albedo,
                                                          const float* __restrict__ discharge, long long n, rgb col,
                                                          float extinction, float scale) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const float cd[3] = {col.r, col.g, col.b};
  if (VEC) {
    const long long i0 = 4 * t;
    if (i0 >= n) return;
    const float4 d = __ldcs(reinterpret_cast<const float4*>(discharge + i0));
    const float dv[4] = {d.x, d.y, d.z, d.w};
    quad12 q = load12(albedo + 3 * i0);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float blend = discharge_blend(dv[j], extinction, scale);
#pragma unroll
      for (int c = 0; c < 3; ++c) q.v[3 * j + c] = blend * cd[c] + (1.0f - blend) * q.v[3 * j + c];
    }
    store12(albedo + 3 * i0, q);
  } else {
    if (t >= n) return;
    const float blend = discharge_blend(__ldg(discharge + t), extinction, scale);
#pragma unroll
    for (int c = 0; c < 3; ++c) albedo[3 * t + c] = blend * cd[c] + (1.0f - blend) * albedo[3 * t + c];
  }
}

inline bool vec_ok(long long n, std::initializer_list<const void*> ptrs) {
  if (n % 4) return false;
  for (const void* p : ptrs)
    if (reinterpret_cast<uintptr_t>(p) & 15) return false;
  return true;
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

int soilb200_albedo_layer(float* albedo, const float* albedo_bedrock, const float* albedo_sediment, const float* layers,
                          int nx, int ny, float scale_sediment, const float shift_sediment[3], void* stream) {
  if (!albedo || !albedo_bedrock || !albedo_sediment || !layers || !shift_sediment || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  const long long n = (long long)nx * ny;
  const rgb sh{shift_sediment[0], shift_sediment[1], shift_sediment[2]};
  cudaStream_t s = (cudaStream_t)stream;
  if (vec_ok(n, {albedo, albedo_bedrock, albedo_sediment, layers}))
    k_albedo_layer<true><<<blocks_for(n / 4, 256), 256, 0, s>>>(albedo, albedo_bedrock, albedo_sediment, (const float2*)layers, n, scale_sediment, sh);
  else
    k_albedo_layer<false><<<blocks_for(n, 256), 256, 0, s>>>(albedo, albedo_bedrock, albedo_sediment, (const float2*)layers, n, scale_sediment, sh);
  count_launch();
  return check_launch();
}

int soilb200_albedo_stratum(float* albedo_bedrock, const float* uplift, const float* layers, int nx, int ny,
                            const float scale[3], const soilb200_param_t* param, const float color_a[3],
                            const float color_b[3], float age, float freq, void* stream) {
  if (!albedo_bedrock || !uplift || !layers || !scale || !param || !color_a || !color_b || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  const long long n = (long long)nx * ny;
  const rgb A{color_a[0], color_a[1], color_a[2]}, B{color_b[0], color_b[1], color_b[2]};
  const float age_uplift = age * param->uplift;  // erosion.cu:809, left to right
  cudaStream_t s = (cudaStream_t)stream;
  if (vec_ok(n, {albedo_bedrock, uplift, layers}))
    k_albedo_stratum<true><<<blocks_for(n / 4, 256), 256, 0, s>>>(albedo_bedrock, uplift, (const float2*)layers, n, age_uplift, scale[2], freq, A, B);
  else
    k_albedo_stratum<false><<<blocks_for(n, 256), 256, 0, s>>>(albedo_bedrock, uplift, (const float2*)layers, n, age_uplift, scale[2], freq, A, B);
  count_launch();
  return check_launch();
}

int soilb200_albedo_discharge(float* albedo, const float* discharge, int nx, int ny, const float color_discharge[3],
                              float extinction, float scale, void* stream) {
  if (!albedo || !discharge || !color_discharge || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  const long long n = (long long)nx * ny;
  const rgb col{color_discharge[0], color_discharge[1], color_discharge[2]};
  cudaStream_t s = (cudaStream_t)stream;
  if (vec_ok(n, {albedo, discharge}))
    k_albedo_discharge<true><<<blocks_for(n / 4, 256), 256, 0, s>>>(albedo, discharge, n, col, extinction, scale);
  else
    k_albedo_discharge<false><<<blocks_for(n, 256), 256, 0, s>>>(albedo, discharge, n, col, extinction, scale);
  count_launch();
  return check_launch();
}

}  // extern "C"
