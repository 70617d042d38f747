// transport.cu - stochastic particle descent: fluvial and debris-flow transport.
//
// Reference: source/soillib/model/path/erosion.cu:29-141 (fluvial particle),
// :245-351 (debris particle).  One particle per thread; a particle is a serial
// chain of <= maxage cell crossings, each a dependent gather of the 5-point
// height stencil + velocity + water height followed by fire-and-forget
// reductions into the flux maps.  The trajectory arithmetic is the reference's,
// operand for operand (no FMA contraction), so a particle visits exactly the
// cells the CPU oracle's particle visits.
//
// Two memory layouts, selected by a policy type:
//  * ref_io   - the reference's tensors as they are (per-op entry points):
//               heights from the interleaved `layers`, one RED per flux float.
//  * fused_io - the fused step's private layout: heights from the merged
//               `height` map (the dense pass writes bedrock + sediment there, the
//               same float add the reference performs per access), and ONE packed
//               flux record per cell and particle system, updated with 16-byte
//               vector reductions (red.global.add.v4.f32): a crossing dirties one
//               32-byte sector instead of five, and tags the 4-cell (one 128-byte
//               line of records) group it touched for the dense pass.
#include "common.cuh"
#include "erosion_math.cuh"
#include <curand_kernel.h>

namespace sb {
namespace {

__device__ __forceinline__ bool oobf(float px, float py, int nx, int ny) {  // erosion_map.cu:29-40
  if (px < 0) return true;
  if (py < 0) return true;
  if (px >= nx) return true;
  if (py >= ny) return true;
  return false;
}

// shape.flatten(ivec2(pos)): truncation toward zero, NaN -> 0 (cvt.rzi), App. A-5 / B-4
__device__ __forceinline__ int flattenf(float px, float py, int ny) { return (int)px * ny + (int)py; }

__device__ __forceinline__ void red_add2(float2* addr, float vx, float vy) {
  atomicAdd(addr, make_float2(vx, vy));  // red.global.add.v2.f32 (sm_90+)
}
__device__ __forceinline__ void red_add4(float* addr, float a, float b, float c, float d) {
  atomicAdd(reinterpret_cast<float4*>(addr), make_float4(a, b, c, d));  // red.global.add.v4.f32 (sm_90+)
}

// ---- layout policies ---------------------------------------------------------------

struct ref_io {
  static constexpr bool kFused = false;
  const float2* layers;
  // fluvial: water, mass, velocity(2), albedo(3); debris: mass, velocity(2), albedo(3)
  float *waterFlux, *massFlux;
  float2* velocityFlux;
  float* albedoFlux;  // may be null
  __device__ __forceinline__ float height(size_t i) const {
    const float2 l = __ldg(layers + i);
    return l.x + l.y;  // erosion_map.cu:99-105
  }
  __device__ __forceinline__ bool albedo() const { return albedoFlux != nullptr; }
  __device__ __forceinline__ void deposit_fluvial(int ind, float w, float m, float vx, float vy, float ar, float ag,
                                                  float ab) const {
    atomicAdd(waterFlux + ind, w);
    atomicAdd(massFlux + ind, m);
    red_add2(velocityFlux + ind, vx, vy);
    if (albedoFlux) {
      atomicAdd(albedoFlux + 3 * (size_t)ind + 0, ar);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 1, ag);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 2, ab);
    }
  }
  __device__ __forceinline__ void deposit_debris(int ind, float m, float vx, float vy, float ar, float ag,
                                                 float ab) const {
    atomicAdd(massFlux + ind, m);
    red_add2(velocityFlux + ind, vx, vy);
    if (albedoFlux) {
      atomicAdd(albedoFlux + 3 * (size_t)ind + 0, ar);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 1, ag);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 2, ab);
    }
  }
};

// record of R = 4 (no albedo) or 8 floats per cell:
//   fluvial {water, mass, vel.x, vel.y | r, g, b, 0}   debris {mass, vel.x, vel.y, 0 | r, g, b, 0}
struct fused_io {
  static constexpr bool kFused = true;
  const float* hmap;
  float* rec;  // this particle system's records
  int R;
  unsigned char* touched;
  __device__ __forceinline__ float height(size_t i) const { return __ldg(hmap + i); }
  __device__ __forceinline__ bool albedo() const { return R == 8; }
  __device__ __forceinline__ void deposit_fluvial(int ind, float w, float m, float vx, float vy, float ar, float ag,
                                                  float ab) const {
    float* r = rec + (size_t)R * ind;
    red_add4(r, w, m, vx, vy);
    if (R == 8) red_add4(r + 4, ar, ag, ab, 0.0f);
    touched[ind >> 2] = 1;
  }
  __device__ __forceinline__ void deposit_debris(int ind, float m, float vx, float vy, float ar, float ag,
                                                 float ab) const {
    float* r = rec + (size_t)R * ind;
    red_add4(r, m, vx, vy, 0.0f);
    if (R == 8) red_add4(r + 4, ar, ag, ab, 0.0f);
    touched[ind >> 2] = 1;
  }
};

// the five heights __glocal (erosion_map.cu:107-159) reads at the cell containing (fx,fy); NaN = outside
struct h5 {
  float c, n0, p0, zn, zp;
};
template <class IO>
__device__ __forceinline__ h5 load5(const IO& io, float fx, float fy, int nx, int ny) {
  const int px = (int)fx, py = (int)fy;
  const size_t i = (size_t)px * ny + py;
  h5 w;
  w.n0 = w.p0 = w.zn = w.zp = CUDART_NAN_F;
  w.c = io.height(i);
  if (px > 0) w.n0 = io.height(i - ny);
  if (px + 1 < nx) w.p0 = io.height(i + ny);
  if (py > 0) w.zn = io.height(i - 1);
  if (py + 1 < ny) w.zp = io.height(i + 1);
  return w;
}
__device__ __forceinline__ f2 grad_of(const h5& w, const scale3 s, float exitSlope) {
  return glocal_h(w.c, w.n0, w.p0, w.zn, w.zp, s, exitSlope);
}

struct particle_ctx {
  curandStateXORWOW_t* rng;
  long long P;
  int nx, ny;
  scale3 scale;
  const float* albedoSource;  // (nx,ny,3) or null
};

// ---- fluvial particle, erosion.cu:46-139 ---------------------------------------------
template <class IO>
__device__ __forceinline__ void fluvial_particle(const long long n, const IO io, const particle_ctx c,
                                                 const float* __restrict__ waterSource,
                                                 const float* __restrict__ waterHeight,
                                                 const float2* __restrict__ velocity, const soilb200_param_t& param) {
  const int nx = c.nx, ny = c.ny;
  const scale3 scale = c.scale;
  const float A = scale.x * scale.y;
  const float Lx = scale.x, Ly = scale.y;
  const float Pp = 1.0f / float(A * (float)((size_t)nx * (size_t)ny));
  const float Q = 1.0f / (Pp * (float)(size_t)c.P);
  const float eps = 1E-12f;

  curandStateXORWOW_t st = c.rng[n];
  float px = 0.5f + curand_uniform(&st) * float(nx - 1);
  float py = 0.5f + curand_uniform(&st) * float(ny - 1);
  c.rng[n] = st;
  int ind = flattenf(px, py, ny);

  const float rho_w = param.densityWater;
  const float tau = param.bedShearWater;
  const float nu = param.viscosityWater;
  const float g = param.gravity;
  const float ks = param.suspensionRateFluvial / 64.0f;
  const float kd = param.depositionRateFluvial * 1.33f;
  const float fD = param.frictionFactor / 8.0f;
  const float alpha = param.fluvialExponent;
  const float R = param.rainfall;

  const float2 vel = __ldg(velocity + ind);
  f2 grad = grad_of(load5(io, px, py, nx, ny), scale, param.exitSlope);
  float sx = -(g * grad.x) + nu * vel.x + param.force[0];
  float sy = -(g * grad.y) + nu * vel.y + param.force[1];
  {
    const float d = sqrtf(len2(Lx * sx, Ly * sy));
    sx = sx / d;
    sy = sy / d;
  }
  if (len2(sx, sy) < eps) return;

  const float v = len2(vel.x, vel.y);
  const float shear = 0.125f * fD * rho_w * v * v;
  const float power = __powf(shear * len2(grad.x, grad.y), alpha);

  const float source_m = Q * ks * power;
  const float source_w = Q * R * __ldg(waterSource + ind);
  const float source_vx = Q * (-(g * grad.x) + nu * vel.x);
  const float source_vy = Q * (-(g * grad.y) + nu * vel.y);
  float source_ax = 0.0f, source_ay = 0.0f, source_az = 0.0f;
  if (io.albedo()) {
    source_ax = source_m * __ldg(c.albedoSource + 3 * (size_t)ind + 0);
    source_ay = source_m * __ldg(c.albedoSource + 3 * (size_t)ind + 1);
    source_az = source_m * __ldg(c.albedoSource + 3 * (size_t)ind + 2);
  }

  float att_w = 1.0f, att_m = 1.0f, att_v = 1.0f;
  const float lenL = len2(Lx, Ly);

  int iter = 0;
  while (!oobf(px, py, nx, ny) && (unsigned long long)(++iter) < param.maxage) {
    const int nind = flattenf(px, py, ny);
    // every gather of this crossing is issued here, before anything is consumed: one memory round
    // trip per crossing instead of three (the reference reads them further down, same addresses)
    const h5 hw = load5(io, px, py, nx, ny);
    const float2 vl = __ldg(velocity + nind);
    const float wh = __ldg(waterHeight + nind);
    if (nind != ind) {
      ind = nind;
      io.deposit_fluvial(ind, att_w * source_w, att_m * source_m, att_v * source_vx, att_v * source_vy,
                         att_m * source_ax, att_m * source_ay, att_m * source_az);
    }

    const float v_norm = len2(sx, sy);
    const float ux = sx / v_norm, uy = sy / v_norm;
    const float v_step = stepsize(px, py, ux, uy);
    const float dL = v_step * lenL;
    const float ds = dL / v_norm;
    if (v_norm < eps) break;

    grad = grad_of(hw, scale, param.exitSlope);
    const float ax = -(g * grad.x) + nu * vl.x + param.force[0];
    const float ay = -(g * grad.y) + nu * vl.y + param.force[1];
    const float c0 = 1.0f / (1.0f + dL * (tau + nu));
    const float c1 = dL / (1.0f + dL * (tau + nu));
    sx = c0 * sx + c1 * ax;
    sy = c0 * sy + c1 * ay;

    const float decay_m = kd;
    const float decay_w = param.evapRate;
    const float decay_v = 0.125f * fD / (eps + wh);
    att_m = att_m * __expf(-ds * decay_m);
    att_w = att_w * __expf(-ds * decay_w);
    att_v = att_v * __expf(-dL * decay_v);
    px += v_step * ux;
    py += v_step * uy;
  }
}

// ---- debris particle, erosion.cu:259-349 -----------------------------------------------
template <class IO>
__device__ __forceinline__ void debris_particle(const long long n, const IO io, const particle_ctx c,
                                                const float2* __restrict__ velocity, const soilb200_param_t& param) {
  const int nx = c.nx, ny = c.ny;
  const scale3 scale = c.scale;
  const float A = scale.x * scale.y;
  const float Lx = scale.x, Ly = scale.y;
  const float Pp = 1.0f / float(A * (float)((size_t)nx * (size_t)ny));
  const float Q = 1.0f / (Pp * (float)(size_t)c.P);
  const float eps = 1E-12f;

  curandStateXORWOW_t st = c.rng[n];
  float px = 0.5f + curand_uniform(&st) * float(nx - 1);
  float py = 0.5f + curand_uniform(&st) * float(ny - 1);
  c.rng[n] = st;
  int ind = flattenf(px, py, ny);

  const float theta = param.critSlopeBedrock;
  const float nu = param.viscosityDebris;
  const float tau = param.bedShearDebris;
  const float g = param.gravity;
  const float kl = param.landslideRateDebris;
  const float kdd = param.depositionRateDebris;
  const float kds = param.suspensionRateDebris;
  const float tau_y = param.yieldStress;

  const float2 vel = __ldg(velocity + ind);
  f2 grad = grad_of(load5(io, px, py, nx, ny), scale, param.exitSlope);
  float sx = -(g * grad.x) + nu * vel.x;
  float sy = -(g * grad.y) + nu * vel.y;
  {
    const float d = sqrtf(len2(Lx * sx, Ly * sy));
    sx = sx / d;
    sy = sy / d;
  }
  if (len2(sx, sy) < eps) return;

  const float excessSlope0 = (len2(grad.x, grad.y) - theta);
  const float suspend = fmaxf(0.0f, kl * excessSlope0);
  const float source_d = Q * suspend;
  const float source_vx = Q * (-g * grad.x + nu * vel.x);
  const float source_vy = Q * (-g * grad.y + nu * vel.y);
  float source_ax = 0.0f, source_ay = 0.0f, source_az = 0.0f;
  if (io.albedo()) {
    source_ax = source_d * __ldg(c.albedoSource + 3 * (size_t)ind + 0);
    source_ay = source_d * __ldg(c.albedoSource + 3 * (size_t)ind + 1);
    source_az = source_d * __ldg(c.albedoSource + 3 * (size_t)ind + 2);
  }

  float att_d = 1.0f, att_v = 1.0f;
  const float lenL = len2(Lx, Ly);

  int iter = 0;
  while (!oobf(px, py, nx, ny) && (unsigned long long)(++iter) < param.maxage) {
    const int nind = flattenf(px, py, ny);
    const h5 hw = load5(io, px, py, nx, ny);  // all gathers of this crossing up front
    const float2 vl = __ldg(velocity + nind);
    if (nind != ind) {
      ind = nind;
      io.deposit_debris(ind, att_d * source_d, att_v * source_vx, att_v * source_vy, att_d * source_ax,
                        att_d * source_ay, att_d * source_az);
    }

    const float v_norm = len2(sx, sy);
    const float ux = sx / v_norm, uy = sy / v_norm;
    const float v_step = stepsize(px, py, ux, uy);
    const float dL = v_step * lenL;
    const float ds = dL / v_norm;
    if (v_norm < eps) break;

    grad = grad_of(hw, scale, param.exitSlope);
    const float debrisHeight = eps + att_d * source_d;
    const float ax = -(g * grad.x) + nu * vl.x;
    const float ay = -(g * grad.y) + nu * vl.y;
    const float decay = nu + tau / debrisHeight;
    const float w = 1.0f / (1.0f + dL * decay);
    sx = w * sx + w * dL * ax;
    sy = w * sy + w * dL * ay;

    const float excessSlope = (len2(grad.x, grad.y) - theta);
    const float excessStress = g * (excessSlope - tau_y / debrisHeight);
    const float shearRate = (excessStress < 0.0f) ? kdd : kds;
    const float decay_d = ds * shearRate * excessStress / v_norm;
    const float decay_v = (nu + tau / debrisHeight);
    att_d = att_d * __expf(decay_d);
    att_v = att_v * __expf(-dL * decay_v);
    px += v_step * ux;
    py += v_step * uy;
  }
}

// ---- kernels ------------------------------------------------------------------------------

constexpr int kParticleBlock = 64;  // small blocks spread the few, long-lived warps over all SMs

__global__ void __launch_bounds__(kParticleBlock) k_transport_fluvial(ref_io io, particle_ctx c,
                                                                      const float* waterSource,
                                                                      const float* waterHeight,
                                                                      const float2* velocity, soilb200_param_t param) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= c.P) return;
  fluvial_particle(n, io, c, waterSource, waterHeight, velocity, param);
}

__global__ void __launch_bounds__(kParticleBlock) k_transport_debris(ref_io io, particle_ctx c, const float2* velocity,
                                                                     soilb200_param_t param) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= c.P) return;
  debris_particle(n, io, c, velocity, param);
}

// fused step: both particle systems in ONE launch (they are independent: different flux records,
// different RNG streams), blocks alternating fluvial / debris so that every SM holds warps of both
// kinds and has twice the warps to hide the serial per-particle latency behind
__global__ void __launch_bounds__(kParticleBlock) k_transport_both(fused_io fio, fused_io dio, particle_ctx fc,
                                                                   particle_ctx dc, const float* waterSource,
                                                                   const float* waterHeight, const float2* velocity,
                                                                   const float2* debrisVel, soilb200_param_t param) {
  const bool is_debris = blockIdx.x & 1;
  const long long n = (long long)(blockIdx.x >> 1) * blockDim.x + threadIdx.x;
  if (n >= fc.P) return;
  if (!is_debris) fluvial_particle(n, fio, fc, waterSource, waterHeight, velocity, param);
  else debris_particle(n, dio, dc, debrisVel, param);
}

}  // namespace

// host launchers used by dense.cu's entry points
int launch_transport_fluvial(float* waterFlux, float* massFlux, float* velocityFlux, float* albedoFlux, void* rng,
                             long long P, const float* layers, const float* rainfall, const float* waterHeight,
                             const float* velocity, const float* albedoSource, int nx, int ny, scale3 s,
                             const soilb200_param_t& p, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const ref_io io{(const float2*)layers, waterFlux, massFlux, (float2*)velocityFlux, albedoFlux};
  const particle_ctx c{(curandStateXORWOW_t*)rng, P, nx, ny, s, albedoSource};
  k_transport_fluvial<<<blocks_for(P, kParticleBlock), kParticleBlock, 0, st>>>(io, c, rainfall, waterHeight,
                                                                               (const float2*)velocity, p);
  count_launch();
  return check_launch();
}

int launch_transport_debris(float* massFlux, float* velocityFlux, float* albedoFlux, void* rng, long long P,
                            const float* layers, const float* velocity, const float* albedoSource, int nx, int ny,
                            scale3 s, const soilb200_param_t& p, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const ref_io io{(const float2*)layers, nullptr, massFlux, (float2*)velocityFlux, albedoFlux};
  const particle_ctx c{(curandStateXORWOW_t*)rng, P, nx, ny, s, albedoSource};
  k_transport_debris<<<blocks_for(P, kParticleBlock), kParticleBlock, 0, st>>>(io, c, (const float2*)velocity, p);
  count_launch();
  return check_launch();
}

// flux = {fluvial records [N*R], debris records [N*R]}, R = 8 with albedo else 4
int launch_transport_both(const float* hmap, float* flux, int R, unsigned char* touched, void* rngF, void* rngD,
                          long long P, const float* rainfall, const float* waterHeight, const float* velocity,
                          const float* debrisVel, const float* albedoSource, int nx, int ny, scale3 s,
                          const soilb200_param_t& p, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const size_t N = (size_t)nx * ny;
  const fused_io fio{hmap, flux, R, touched};
  const fused_io dio{hmap, flux + N * R, R, touched};
  const particle_ctx fc{(curandStateXORWOW_t*)rngF, P, nx, ny, s, albedoSource};
  const particle_ctx dc{(curandStateXORWOW_t*)rngD, P, nx, ny, s, albedoSource};
  k_transport_both<<<2 * blocks_for(P, kParticleBlock), kParticleBlock, 0, st>>>(
      fio, dio, fc, dc, rainfall, waterHeight, (const float2*)velocity, (const float2*)debrisVel, p);
  count_launch();
  return check_launch();
}

}  // namespace sb
