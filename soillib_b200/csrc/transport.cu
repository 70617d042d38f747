// transport.cu - stochastic particle descent: fluvial and debris-flow transport.
//
// Reference: source/soillib/model/path/erosion.cu:29-141 (fluvial particle),
// :245-351 (debris particle).  One particle per thread; a particle is a serial
// chain of <= maxage cell crossings, each a dependent gather of the 5-point
// `layers` stencil + velocity + water height followed by fire-and-forget
// reductions into the flux maps.  The trajectory arithmetic is the reference's,
// operand for operand (no FMA contraction), so a particle visits exactly the
// cells the CPU oracle's particle visits.
//
// B200 notes: the gathers go through the read-only path and stay in L1 across
// consecutive crossings (a crossing moves one cell, i.e. mostly within the same
// 128-byte lines); the next rows along the direction of travel are prefetched
// into L2/L1 one crossing ahead; flux updates are RED (no return value), with
// the two velocity components fused into one 8-byte vector reduction.  With
// MARK, every crossing also tags the 32-cell segment it touched so that the
// fused dense pass can skip flux traffic for untouched segments.
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

__device__ __forceinline__ f2 grad_at(const float2* __restrict__ layers, float px, float py, int nx, int ny,
                                      const scale3 s, float exitSlope) {
  return glocal_mem(layers, (int)px, (int)py, nx, ny, s, exitSlope);
}

__device__ __forceinline__ void red_add2(float2* addr, float vx, float vy) {
#if __CUDA_ARCH__ >= 900
  atomicAdd(addr, make_float2(vx, vy));  // red.global.add.v2.f32
#else
  atomicAdd(&addr->x, vx);
  atomicAdd(&addr->y, vy);
#endif
}

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// rows the particle will need after its next crossing(s), one step ahead of the dependent chain
__device__ __forceinline__ void prefetch_ahead(const float2* layers, const float2* vel, const float* wh, float px,
                                               float py, float ux, float uy, int nx, int ny) {
  const int ix = (int)px, iy = (int)py;
  const int sx = ux > 0.0f ? 2 : -2;
  const int tx = ix + sx;
  if (tx >= 0 && tx < nx && iy >= 0 && iy < ny) {
    const size_t i = (size_t)tx * ny + iy;
    prefetch_l1(layers + i);
    const size_t j = (size_t)(tx - sx / 2) * ny + iy;
    prefetch_l1(vel + j);
    if (wh) prefetch_l1(wh + j);
  }
  const int sy = uy > 0.0f ? 16 : -16;  // next 128-byte line of float2 along the row
  const int ty = iy + sy;
  if (ix >= 0 && ix < nx && ty >= 0 && ty < ny) prefetch_l1(layers + (size_t)ix * ny + ty);
}

struct fluvial_args {
  float *waterFlux, *massFlux;
  float2* velocityFlux;
  float* albedoFlux;
  curandStateXORWOW_t* rng;
  const float2* layers;
  const float *waterSource, *waterHeight;
  const float2* velocity;
  const float* albedoSource;
};
struct debris_args {
  float* massFlux;
  float2* velocityFlux;
  float* albedoFlux;
  curandStateXORWOW_t* rng;
  const float2* layers;
  const float2* velocity;
  const float* albedoSource;
};

template <bool MARK>
__device__ __forceinline__ void fluvial_particle(const long long n,
    float* __restrict__ waterFlux, float* __restrict__ massFlux, float2* __restrict__ velocityFlux,
    float* __restrict__ albedoFlux, curandStateXORWOW_t* __restrict__ rng, long long P,
    const float2* __restrict__ layers, const float* __restrict__ waterSource, const float* __restrict__ waterHeight,
    const float2* __restrict__ velocity, const float* __restrict__ albedoSource, int nx, int ny, scale3 scale,
    const soilb200_param_t& param, unsigned char* __restrict__ touched) {
  const float A = scale.x * scale.y;
  const float Lx = scale.x, Ly = scale.y;
  const float Pp = 1.0f / float(A * (float)((size_t)nx * (size_t)ny));
  const float Q = 1.0f / (Pp * (float)(size_t)P);
  const float eps = 1E-12f;

  curandStateXORWOW_t st = rng[n];
  float px = 0.5f + curand_uniform(&st) * float(nx - 1);
  float py = 0.5f + curand_uniform(&st) * float(ny - 1);
  rng[n] = st;
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
  f2 grad = grad_at(layers, px, py, nx, ny, scale, param.exitSlope);
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
  if (albedoFlux) {
    source_ax = source_m * __ldg(albedoSource + 3 * (size_t)ind + 0);
    source_ay = source_m * __ldg(albedoSource + 3 * (size_t)ind + 1);
    source_az = source_m * __ldg(albedoSource + 3 * (size_t)ind + 2);
  }

  float att_w = 1.0f, att_m = 1.0f, att_v = 1.0f;
  const float lenL = len2(Lx, Ly);

  int iter = 0;
  while (!oobf(px, py, nx, ny) && (unsigned long long)(++iter) < param.maxage) {
    const int nind = flattenf(px, py, ny);
    if (nind != ind) {
      ind = nind;
      atomicAdd(waterFlux + ind, att_w * source_w);
      atomicAdd(massFlux + ind, att_m * source_m);
      red_add2(velocityFlux + ind, att_v * source_vx, att_v * source_vy);
      if (albedoFlux) {
        atomicAdd(albedoFlux + 3 * (size_t)ind + 0, att_m * source_ax);
        atomicAdd(albedoFlux + 3 * (size_t)ind + 1, att_m * source_ay);
        atomicAdd(albedoFlux + 3 * (size_t)ind + 2, att_m * source_az);
      }
      if (MARK) touched[ind >> 5] = 1;
    }

    const float v_norm = len2(sx, sy);
    const float ux = sx / v_norm, uy = sy / v_norm;
    const float v_step = stepsize(px, py, ux, uy);
    const float dL = v_step * lenL;
    const float ds = dL / v_norm;
    if (v_norm < eps) break;

    prefetch_ahead(layers, velocity, waterHeight, px, py, ux, uy, nx, ny);

    grad = grad_at(layers, px, py, nx, ny, scale, param.exitSlope);
    const float2 vl = __ldg(velocity + ind);
    const float ax = -(g * grad.x) + nu * vl.x + param.force[0];
    const float ay = -(g * grad.y) + nu * vl.y + param.force[1];
    const float c0 = 1.0f / (1.0f + dL * (tau + nu));
    const float c1 = dL / (1.0f + dL * (tau + nu));
    sx = c0 * sx + c1 * ax;
    sy = c0 * sy + c1 * ay;

    const float decay_m = kd;
    const float decay_w = param.evapRate;
    const float decay_v = 0.125f * fD / (eps + __ldg(waterHeight + ind));
    att_m = att_m * __expf(-ds * decay_m);
    att_w = att_w * __expf(-ds * decay_w);
    att_v = att_v * __expf(-dL * decay_v);
    px += v_step * ux;
    py += v_step * uy;
  }
}

template <bool MARK>
__device__ __forceinline__ void debris_particle(const long long n,
    float* __restrict__ massFlux, float2* __restrict__ velocityFlux, float* __restrict__ albedoFlux,
    curandStateXORWOW_t* __restrict__ rng, long long P, const float2* __restrict__ layers,
    const float2* __restrict__ velocity, const float* __restrict__ albedoSource, int nx, int ny, scale3 scale,
    const soilb200_param_t& param, unsigned char* __restrict__ touched) {
  const float A = scale.x * scale.y;
  const float Lx = scale.x, Ly = scale.y;
  const float Pp = 1.0f / float(A * (float)((size_t)nx * (size_t)ny));
  const float Q = 1.0f / (Pp * (float)(size_t)P);
  const float eps = 1E-12f;

  curandStateXORWOW_t st = rng[n];
  float px = 0.5f + curand_uniform(&st) * float(nx - 1);
  float py = 0.5f + curand_uniform(&st) * float(ny - 1);
  rng[n] = st;
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
  f2 grad = grad_at(layers, px, py, nx, ny, scale, param.exitSlope);
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
  if (albedoFlux) {
    source_ax = source_d * __ldg(albedoSource + 3 * (size_t)ind + 0);
    source_ay = source_d * __ldg(albedoSource + 3 * (size_t)ind + 1);
    source_az = source_d * __ldg(albedoSource + 3 * (size_t)ind + 2);
  }

  float att_d = 1.0f, att_v = 1.0f;
  const float lenL = len2(Lx, Ly);

  int iter = 0;
  while (!oobf(px, py, nx, ny) && (unsigned long long)(++iter) < param.maxage) {
    const int nind = flattenf(px, py, ny);
    if (nind != ind) {
      ind = nind;
      atomicAdd(massFlux + ind, att_d * source_d);
      red_add2(velocityFlux + ind, att_v * source_vx, att_v * source_vy);
      if (albedoFlux) {
        atomicAdd(albedoFlux + 3 * (size_t)ind + 0, att_d * source_ax);
        atomicAdd(albedoFlux + 3 * (size_t)ind + 1, att_d * source_ay);
        atomicAdd(albedoFlux + 3 * (size_t)ind + 2, att_d * source_az);
      }
      if (MARK) touched[ind >> 5] = 1;
    }

    const float v_norm = len2(sx, sy);
    const float ux = sx / v_norm, uy = sy / v_norm;
    const float v_step = stepsize(px, py, ux, uy);
    const float dL = v_step * lenL;
    const float ds = dL / v_norm;
    if (v_norm < eps) break;

    prefetch_ahead(layers, velocity, nullptr, px, py, ux, uy, nx, ny);

    grad = grad_at(layers, px, py, nx, ny, scale, param.exitSlope);
    const float debrisHeight = eps + att_d * source_d;
    const float2 vl = __ldg(velocity + ind);
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

constexpr int kParticleBlock = 64;  // small blocks spread the few, long-lived warps over all SMs

template <bool MARK>
__global__ void __launch_bounds__(kParticleBlock) k_transport_fluvial(fluvial_args a, long long P, int nx, int ny,
                                                                      scale3 scale, soilb200_param_t param,
                                                                      unsigned char* touched) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= P) return;
  fluvial_particle<MARK>(n, a.waterFlux, a.massFlux, a.velocityFlux, a.albedoFlux, a.rng, P, a.layers, a.waterSource,
                         a.waterHeight, a.velocity, a.albedoSource, nx, ny, scale, param, touched);
}

template <bool MARK>
__global__ void __launch_bounds__(kParticleBlock) k_transport_debris(debris_args a, long long P, int nx, int ny,
                                                                     scale3 scale, soilb200_param_t param,
                                                                     unsigned char* touched) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= P) return;
  debris_particle<MARK>(n, a.massFlux, a.velocityFlux, a.albedoFlux, a.rng, P, a.layers, a.velocity, a.albedoSource, nx,
                        ny, scale, param, touched);
}

// fused step: both particle systems in ONE launch (they are independent: different flux maps, different
// RNG streams), blocks alternating fluvial / debris so that every SM holds warps of both kinds and has twice
// the warps to hide the serial per-particle latency behind
template <bool MARK>
__global__ void __launch_bounds__(kParticleBlock) k_transport_both(fluvial_args f, debris_args d, long long P, int nx,
                                                                   int ny, scale3 scale, soilb200_param_t param,
                                                                   unsigned char* touched) {
  const bool is_debris = blockIdx.x & 1;
  const long long n = (long long)(blockIdx.x >> 1) * blockDim.x + threadIdx.x;
  if (n >= P) return;
  if (!is_debris)
    fluvial_particle<MARK>(n, f.waterFlux, f.massFlux, f.velocityFlux, f.albedoFlux, f.rng, P, f.layers, f.waterSource,
                           f.waterHeight, f.velocity, f.albedoSource, nx, ny, scale, param, touched);
  else
    debris_particle<MARK>(n, d.massFlux, d.velocityFlux, d.albedoFlux, d.rng, P, d.layers, d.velocity, d.albedoSource,
                          nx, ny, scale, param, touched);
}

}  // namespace

// host launchers used by dense.cu's entry points
int launch_transport_fluvial(float* waterFlux, float* massFlux, float* velocityFlux, float* albedoFlux, void* rng,
                             long long P, const float* layers, const float* rainfall, const float* waterHeight,
                             const float* velocity, const float* albedoSource, int nx, int ny, scale3 s,
                             const soilb200_param_t& p, unsigned char* touched, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const fluvial_args a{waterFlux, massFlux, (float2*)velocityFlux, albedoFlux, (curandStateXORWOW_t*)rng,
                       (const float2*)layers, rainfall, waterHeight, (const float2*)velocity, albedoSource};
  const unsigned nb = blocks_for(P, kParticleBlock);
  if (touched) k_transport_fluvial<true><<<nb, kParticleBlock, 0, st>>>(a, P, nx, ny, s, p, touched);
  else k_transport_fluvial<false><<<nb, kParticleBlock, 0, st>>>(a, P, nx, ny, s, p, nullptr);
  count_launch();
  return check_launch();
}

int launch_transport_debris(float* massFlux, float* velocityFlux, float* albedoFlux, void* rng, long long P,
                            const float* layers, const float* velocity, const float* albedoSource, int nx, int ny,
                            scale3 s, const soilb200_param_t& p, unsigned char* touched, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const debris_args a{massFlux, (float2*)velocityFlux, albedoFlux, (curandStateXORWOW_t*)rng, (const float2*)layers,
                      (const float2*)velocity, albedoSource};
  const unsigned nb = blocks_for(P, kParticleBlock);
  if (touched) k_transport_debris<true><<<nb, kParticleBlock, 0, st>>>(a, P, nx, ny, s, p, touched);
  else k_transport_debris<false><<<nb, kParticleBlock, 0, st>>>(a, P, nx, ny, s, p, nullptr);
  count_launch();
  return check_launch();
}

int launch_transport_both(float* waterFlux, float* massFlux, float* velocityFlux, float* albedoFluxF, void* rngF,
                          const float* rainfall, const float* waterHeight, const float* velocity, float* debrisFlux,
                          float* debrisVelFlux, float* albedoFluxD, void* rngD, const float* debrisVel, long long P,
                          const float* layers, const float* albedoSource, int nx, int ny, scale3 s,
                          const soilb200_param_t& p, unsigned char* touched, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const fluvial_args f{waterFlux, massFlux, (float2*)velocityFlux, albedoFluxF, (curandStateXORWOW_t*)rngF,
                       (const float2*)layers, rainfall, waterHeight, (const float2*)velocity, albedoSource};
  const debris_args d{debrisFlux, (float2*)debrisVelFlux, albedoFluxD, (curandStateXORWOW_t*)rngD,
                      (const float2*)layers, (const float2*)debrisVel, albedoSource};
  const unsigned nb = 2 * blocks_for(P, kParticleBlock);
  k_transport_both<true><<<nb, kParticleBlock, 0, st>>>(f, d, P, nx, ny, s, p, touched);
  count_launch();
  return check_launch();
}

}  // namespace sb
