// erosion_math.cuh - per-cell / per-particle arithmetic of the erosion model,
// written once and used by the per-op kernels and by the fused step so that
// both produce the same bits.
//
// Reference: source/soillib/model/path/erosion_map.cu:29-171 (helpers, one-sided
// downhill gradient), erosion.cu:143-187 / :353-393 (normalise), :471-572
// (transfer), :641-708 (creep).  Every expression keeps the reference's
// operand order; the library is built with -fmad=false.
#pragma once
#include "common.cuh"

namespace sb {

struct f2 {
  float x, y;
};
struct f3 {
  float x, y, z;
};

// Sub-expressions of the reference's kernels that depend only on param and scale.  Every thread of
// the reference recomputes them; here the host evaluates them once with the same binary32 operations
// in the same order (host code is built with -ffp-contract=off), so the per-cell results are unchanged.
// a / b, IEEE round-to-nearest like the plain division, but a numerator that is exactly zero does not take the
// division's slow path (a subroutine that runs one lane at a time and first waits for every outstanding load of
// the warp): over a normal divisor the quotient is a zero whose sign is sign(a) xor sign(b).  In the particle loop
// zero numerators are common - a particle that stands exactly on a cell edge has a zero distance to it, hence a
// zero step, a zero path length and a zero time of flight (measured: 2 % of all crossings, i.e. one lane in 45 % of
// the warp iterations, four such divisions each); in the dense pass a cell no particle touched has zero mass fluxes
// and half the map has zero uplift.
__device__ __forceinline__ float div_zn(float a, float b) {
  const float ab = fabsf(b);
  const bool z = (a == 0.0f) && (ab >= 1.17549435e-38f) && (ab <= 3.402823466e+38f);
  const float q = (z ? 1.0f : a) / b;
  return z ? __int_as_float((__float_as_int(a) ^ __float_as_int(b)) & 0x80000000) : q;
}

struct derived_t {
  float kfs;          // suspensionRateFluvial / 64.0f            erosion.cu:478
  float kfd;          // depositionRateFluvial * 1.33f            erosion.cu:479
  float c_shear;      // 0.125f * (frictionFactor / 8.0f) * densityWater   erosion.cu:480,499
  float L;            // length(vec2(scale.x, scale.y))           erosion.cu:493
  float negQuarterL;  // -0.25f * L                               erosion.cu:527
  float capDeposit;   // 0.25f * L * 0.3f                         erosion.cu:528
  float A;            // scale.x * scale.y                        erosion.cu:163
  float norm;         // |1 * scale.y| + |0 * scale.x|            erosion.cu:165-166
  float norm_r;       // its reciprocal and kind (common.cuh, cdiv), when the scale was built by make_scale
  int norm_k;
};
inline derived_t derive(const soilb200_param_t& p, const scale3 s) {
  derived_t d;
  d.kfs = p.suspensionRateFluvial / 64.0f;
  d.kfd = p.depositionRateFluvial * 1.33f;
  const float fD = p.frictionFactor / 8.0f;
  d.c_shear = 0.125f * fD * p.densityWater;
  d.L = sqrtf(s.x * s.x + s.y * s.y);
  d.negQuarterL = -0.25f * d.L;
  d.capDeposit = 0.25f * d.L * 0.3f;
  d.A = s.x * s.y;
  d.norm = fabsf(1.0f * s.y) + fabsf(0.0f * s.x);
  d.norm_r = 0.0f;
  d.norm_k = (s.kx | s.ky | s.kz) ? divisor_kind(d.norm, &d.norm_r) : 0;
  return d;
}

// erosion_map.cu:131-157 given the five heights; a NaN neighbour means "outside the map"
__device__ __forceinline__ f2 glocal_h(float h, float hn0, float hp0, float h0n, float h0p, const scale3 s,
                                       float exitSlope) {
  float gxn = cdiv((h - hn0) * s.z, s.x, s.rx, s.kx);
  if (gxn != gxn) gxn = exitSlope; else gxn = fmaxf(gxn, 0.0f);
  float gyn = cdiv((h - h0n) * s.z, s.y, s.ry, s.ky);
  if (gyn != gyn) gyn = exitSlope; else gyn = fmaxf(gyn, 0.0f);
  float gxp = cdiv((hp0 - h) * s.z, s.x, s.rx, s.kx);
  if (gxp != gxp) gxp = -exitSlope; else gxp = fminf(gxp, 0.0f);
  float gyp = cdiv((h0p - h) * s.z, s.y, s.ry, s.ky);
  if (gyp != gyp) gyp = -exitSlope; else gyp = fminf(gyp, 0.0f);
  float gx = 0.0f;
  if (fabsf(gxn) > fabsf(gx)) gx = gxn;
  if (fabsf(gxp) > fabsf(gx)) gx = gxp;
  float gy = 0.0f;
  if (fabsf(gyn) > fabsf(gy)) gy = gyn;
  if (fabsf(gyp) > fabsf(gy)) gy = gyp;
  return {gx, gy};
}

// __glocal straight from memory (particles; per-op kernels)
__device__ __forceinline__ f2 glocal_mem(const float2* __restrict__ layers, int px, int py, int nx, int ny,
                                         const scale3 s, float exitSlope) {
  const size_t i = (size_t)px * ny + py;
  const float2 c = __ldg(layers + i);
  const float h = c.x + c.y;
  float hn0 = CUDART_NAN_F, hp0 = CUDART_NAN_F, h0n = CUDART_NAN_F, h0p = CUDART_NAN_F;
  if (px > 0) { const float2 l = __ldg(layers + i - ny); hn0 = l.x + l.y; }
  if (px + 1 < nx) { const float2 l = __ldg(layers + i + ny); hp0 = l.x + l.y; }
  if (py > 0) { const float2 l = __ldg(layers + i - 1); h0n = l.x + l.y; }
  if (py + 1 < ny) { const float2 l = __ldg(layers + i + 1); h0p = l.x + l.y; }
  return glocal_h(h, hn0, hp0, h0n, h0p, s, exitSlope);
}

// erosion_map.cu:56-78
__device__ __forceinline__ float stepsize(float px, float py, float dx, float dy) {
  const float tmax = CUDART_SQRT_TWO_F;
  const float x_neg = floorf(px);
  const float y_neg = floorf(py);
  const float x_pos = 1.0f + x_neg;
  const float y_pos = 1.0f + y_neg;
  const float tx_neg = div_zn(x_neg - px, dx);
  const float tx_pos = (x_pos - px) / dx;
  const float tx = fminf(fmaxf(tx_neg, tx_pos), tmax);
  const float ty_neg = div_zn(y_neg - py, dy);
  const float ty_pos = (y_pos - py) / dy;
  const float ty = fminf(fmaxf(ty_neg, ty_pos), tmax);
  return 0.5f * (tx + ty);
}

// x / norm (erosion.cu:165-185): a launch constant; zero numerators (the mass fluxes of a cell no particle touched) skip
// the division's slow path either way
__device__ __forceinline__ float ndiv(float x, const derived_t& dv) {
  return dv.norm_k ? cdiv(x, dv.norm, dv.norm_r, dv.norm_k) : div_zn(x, dv.norm);
}
__device__ __forceinline__ float zdiv(float x, const scale3& s) {  // x / scale.z
  return s.kz ? cdiv(x, s.z, s.rz, s.kz) : div_zn(x, s.z);
}

// erosion.cu:160-185 (fluvial) - the caller supplies grad and the raw fluxes
struct fluvial_norm {
  float waterHeight, mass;
  f2 velocity;
};
__device__ __forceinline__ fluvial_norm normalize_fluvial(f2 grad, float rain, float waterFlux, float massFlux,
                                                          f2 velFlux, const derived_t& dv, const soilb200_param_t& p) {
  const float A = dv.A;
  // norm: velocity hard-wired to (1,0), erosion.cu:165-166 (derive)
  const float source_w = p.rainfall * rain;
  const f2 source_v = {-p.gravity * grad.x + p.force[0], -p.gravity * grad.y + p.force[1]};
  const float source_m = 0.0f;
  fluvial_norm o;
  o.waterHeight = ndiv(A * source_w + waterFlux, dv);
  o.mass = ndiv(A * source_m + massFlux, dv);
  o.velocity.x = ndiv(A * source_v.x + velFlux.x, dv);
  o.velocity.y = ndiv(A * source_v.y + velFlux.y, dv);
  return o;
}

// erosion.cu:367-391 (debris)
struct debris_norm {
  float mass;
  f2 velocity;
};
__device__ __forceinline__ debris_norm normalize_debris(f2 grad, float massFlux, f2 velFlux, const derived_t& dv,
                                                        const soilb200_param_t& p) {
  const float A = dv.A;
  const f2 source_v = {-p.gravity * grad.x, -p.gravity * grad.y};
  const float source_d = 0.0f;
  debris_norm o;
  o.mass = ndiv(A * source_d + massFlux, dv);
  o.velocity.x = ndiv(A * source_v.x + velFlux.x, dv);
  o.velocity.y = ndiv(A * source_v.y + velFlux.y, dv);
  return o;
}

// erosion.cu:181-185 / :387-391: accumulated colour -> colour (vec3 length taken on .xy, App. B-5)
__device__ __forceinline__ f3 albedo_norm(float m, f3 a, f3 albedoSource) {
  if (m > 0.0f && len2(a.x, a.y) > 0.0f) return {a.x / m, a.y / m, a.z / m};
  return albedoSource;
}

// erosion.cu:476-547: returns the residual `transfer` that the albedo branch tests (App. B-11)
__device__ __forceinline__ float transfer_cell(f2 grad, f2 layer, f2& delta, float upliftBase, float massHeight,
                                               f2 speed, float debrisHeight, const scale3 s, const derived_t& dv,
                                               const soilb200_param_t& p) {
  const float dt = p.timeStep;
  const float ku = p.uplift;
  const float kfs = dv.kfs;
  const float kfd = dv.kfd;
  const float alpha = p.fluvialExponent;
  const float g = p.gravity;
  const float tau_y = p.yieldStress;
  const float kds = p.suspensionRateDebris;
  const float kdd = p.depositionRateDebris;
  const float kL = p.landslideRateDebris;

  const float slope = len2(grad.x, grad.y);

  const float v = len2(speed.x, speed.y);
  const float shear = dv.c_shear * v * v;
  const float power = __powf(shear * slope, alpha);
  const float suspend = kfs * power;

  const float deposit = kfd * massHeight;
  const float uplift = ku * upliftBase;

  const float excessSlope = (slope - p.critSlopeBedrock);
  const float shearLandslide = fmaxf(0.0f, kL * excessSlope);
  const float shearYield = g * (debrisHeight * excessSlope - tau_y);
  const float suspendDebris = shearLandslide + kds * fmaxf(0.0f, shearYield);
  const float depositDebris = fminf(debrisHeight, fmaxf(0.0f, -kdd * shearYield));

  float transfer = dt * (deposit - suspend + depositDebris - suspendDebris);
  transfer = fmaxf(transfer, dv.negQuarterL * slope);
  transfer = fminf(transfer, dv.capDeposit);

  delta.x += zdiv(dt * uplift, s);
  delta.y += fmaxf(0.0f, zdiv(transfer, s));
  if (transfer < 0.0f) {
    const float limited = fmaxf(-layer.y * s.z, transfer);
    delta.y += zdiv(limited, s);
    transfer -= limited;
    delta.x += zdiv(transfer, s);
  }
  return transfer;
}

__device__ __forceinline__ f3 mmin3(float f, f3 v) { return {fminf(v.x, f), fminf(v.y, f), fminf(v.z, f)}; }

// erosion.cu:553-572; returns true when albedo_surface must be rewritten with `out`
template <class LoadBedrock, class LoadSurface>
__device__ __forceinline__ bool albedo_mix(f2 layer, float transfer, float massHeight, float debrisHeight,
                                           f3 albedoFluvial, f3 albedoDebris, const scale3 s, LoadBedrock bedrock,
                                           LoadSurface surface, f3& out) {
  const float eps = 1E-12f;
  const float totalHeight = massHeight + debrisHeight;
  const float mixDepth = 1.0f;
  if (layer.y == 0.0f) {
    out = bedrock();
    return true;
  }
  if (totalHeight > 0.0f && transfer > eps) {
    const float wMass = fminf(massHeight / totalHeight, 1.0f);
    const f3 colorTransport = mmin3(1.0f, {wMass * albedoFluvial.x + (1.0f - wMass) * albedoDebris.x,
                                           wMass * albedoFluvial.y + (1.0f - wMass) * albedoDebris.y,
                                           wMass * albedoFluvial.z + (1.0f - wMass) * albedoDebris.z});
    const f3 colorSurface = mmin3(1.0f, surface());
    const float wSurf = fminf(mixDepth, layer.y * s.z);
    const float wTrsp = fmaxf(eps, transfer);
    const float w = fminf(wTrsp / (wTrsp + wSurf), 1.0f);
    out = {w * colorTransport.x + (1.0f - w) * colorSurface.x, w * colorTransport.y + (1.0f - w) * colorSurface.y,
           w * colorTransport.z + (1.0f - w) * colorSurface.z};
    return true;
  }
  return false;
}

// erosion.cu:675-680 applied to one neighbour q of cell c (erosion.cu:684-706): the higher cell is the
// "top" and gives; `t += give` when the neighbour is higher, `t -= give` otherwise.  Branch-free select of
// top/bottom (the reference evaluates one of two calls), same operands, same result.
__device__ __forceinline__ float creep_pair(f2 lc, f2 lq, float hc, float hq, float dx, float critSlope, float sz) {
  const bool up = hq > hc;
  const float ht = up ? hq : hc;
  const float hb = up ? hc : hq;
  const float sed_top = up ? lq.y : lc.y;
  const float tmax = 0.5f * ((ht - hb) - critSlope * dx);
  const float give = fmaxf(0.0f, fminf(sed_top * sz, tmax));
  return up ? give : -give;
}

// erosion.cu:660-708; neighbours already clamped to self at the map edge (:654-658)
__device__ __forceinline__ float creep_cell(f2 l00, f2 ln0, f2 lp0, f2 l0n, f2 l0p, const scale3 s, float critSlope) {
  const float h00 = (l00.x + l00.y) * s.z;
  const float hn0 = (ln0.x + ln0.y) * s.z;
  const float hp0 = (lp0.x + lp0.y) * s.z;
  const float h0n = (l0n.x + l0n.y) * s.z;
  const float h0p = (l0p.x + l0p.y) * s.z;
  float t = 0.0f;
  t += creep_pair(l00, lp0, h00, hp0, s.x, critSlope, s.z);
  t += creep_pair(l00, ln0, h00, hn0, s.x, critSlope, s.z);
  t += creep_pair(l00, l0p, h00, h0p, s.y, critSlope, s.z);
  t += creep_pair(l00, l0n, h00, h0n, s.y, critSlope, s.z);
  return zdiv(0.25f * t, s);  // added to delta.y
}

}  // namespace sb
