// dense.cu - the full-grid passes of the erosion step and the fused step.
//
// Reference: source/soillib/model/path/erosion.cu:143-187 (__normalize_fluvial),
// :353-393 (__normalize_debris), :453-611 (__transfer), :633-727 (__mass_creep),
// :733-757 (__layer_merge).  The reference runs them as separate 512-thread
// launches with a zeroing memset per flux map and an add for the apply step:
// ~372 B/cell-step of compulsory traffic.
//
// The per-op entry points below keep that granularity (one kernel each, the
// reference's in-place contracts).  soilb200_erode_step is the B200 path: two
// particle kernels and ONE dense kernel that, per cell, computes the downhill
// gradient once, normalises both transports, evaluates transfer and creep
// against the pre-update layers, applies the result into the ping-pong layer
// buffer and writes the merged height - reading every state map once and
// writing it once.  Particle fluxes live in packed per-cell records (one
// 32-byte sector per cell and particle system) that are only read for cells
// a particle actually touched this step (the particle kernel sets a tag bit per cell);
// the records of the previous step are re-zeroed with pure stores (ping-pong),
// because read-then-zero of the same lines runs ~5x below copy bandwidth on B200.
#include "common.cuh"
#include "erosion_math.cuh"
#include <vector>
#include <cstring>
#include <cstdlib>
#include <cuda.h>

namespace sb {

int launch_transport_fluvial(float*, float*, float*, float*, void*, long long, const float*, const float*,
                             const float*, const float*, const float*, int, int, scale3, const soilb200_param_t&,
                             cudaStream_t);
int launch_transport_debris(float*, float*, float*, void*, long long, const float*, const float*, const float*, int,
                            int, scale3, const soilb200_param_t&, cudaStream_t);

int launch_transport_both(const float*, float*, int, unsigned char*, void*, void*, long long, const float*,
                          const float*, const float*, const float*, const float*, int, int, scale3,
                          const soilb200_param_t&, const strip_host&, void*, const void*, cudaStream_t);
int launch_spawn(void*, void*, long long, int, const strip_host&, void*, cudaStream_t);
int launch_transport_resume(const float*, float*, int, unsigned char*, const float*, int, long long, const float*,
                            const float*, const float*, int, int, scale3, const soilb200_param_t&, const strip_host&,
                            const void*, cudaStream_t);

namespace {

__device__ __forceinline__ f3 ld3(const float* p, size_t n) { return {p[3 * n], p[3 * n + 1], p[3 * n + 2]}; }
__device__ __forceinline__ void st3(float* p, size_t n, f3 v) { p[3 * n] = v.x; p[3 * n + 1] = v.y; p[3 * n + 2] = v.z; }

// ---- per-op kernels (reference granularity) -----------------------------------

__global__ void __launch_bounds__(256) k_normalize_fluvial(
    const float* __restrict__ waterFlux, const float* __restrict__ massFlux, const float2* __restrict__ velocityFlux,
    float* __restrict__ albedoFlux, const float2* __restrict__ layers, const float* __restrict__ waterSource,
    float* __restrict__ waterHeight, float* __restrict__ mass, float2* __restrict__ velocity,
    const float* __restrict__ albedoSource, int nx, int ny, scale3 s, derived_t dv, soilb200_param_t p) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  const f2 grad = glocal_mem(layers, x, y, nx, ny, s, p.exitSlope);
  const float m = massFlux[n];
  const float2 vf = velocityFlux[n];
  const fluvial_norm o = normalize_fluvial(grad, waterSource[n], waterFlux[n], m, {vf.x, vf.y}, dv, p);
  waterHeight[n] = o.waterHeight;
  mass[n] = o.mass;
  velocity[n] = make_float2(o.velocity.x, o.velocity.y);
  if (albedoFlux) st3(albedoFlux, n, albedo_norm(m, ld3(albedoFlux, n), ld3(albedoSource, n)));
}

__global__ void __launch_bounds__(256) k_normalize_debris(
    const float* __restrict__ massFlux, const float2* __restrict__ velocityFlux, float* __restrict__ albedoFlux,
    const float2* __restrict__ layers, float* __restrict__ mass, float2* __restrict__ velocity,
    const float* __restrict__ albedoSource, int nx, int ny, scale3 s, derived_t dv, soilb200_param_t p) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  const f2 grad = glocal_mem(layers, x, y, nx, ny, s, p.exitSlope);
  const float m = massFlux[n];
  const float2 vf = velocityFlux[n];
  const debris_norm o = normalize_debris(grad, m, {vf.x, vf.y}, dv, p);
  mass[n] = o.mass;
  velocity[n] = make_float2(o.velocity.x, o.velocity.y);
  if (albedoFlux) st3(albedoFlux, n, albedo_norm(m, ld3(albedoFlux, n), ld3(albedoSource, n)));
}

__global__ void __launch_bounds__(256) k_transfer(
    float2* __restrict__ deltas, const float2* __restrict__ layers, const float* __restrict__ upliftBase,
    const float* __restrict__ mass, const float2* __restrict__ velocityFluvial, const float* __restrict__ debris,
    const float* __restrict__ albedo_bedrock, const float* __restrict__ albedoFluvial,
    const float* __restrict__ albedoDebris, float* __restrict__ albedo_surface, int nx, int ny, scale3 s,
    derived_t dv, soilb200_param_t p) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  const f2 grad = glocal_mem(layers, x, y, nx, ny, s, p.exitSlope);
  const float2 l = layers[n];
  const float2 d0 = deltas[n];
  const float2 sp = velocityFluvial[n];
  f2 delta = {d0.x, d0.y};
  const float massHeight = mass[n], debrisHeight = debris[n];
  const float transfer = transfer_cell(grad, {l.x, l.y}, delta, upliftBase[n], massHeight, {sp.x, sp.y}, debrisHeight, s, dv, p);
  deltas[n] = make_float2(delta.x, delta.y);
  if (!albedo_surface) return;
  f3 out;
  const size_t i = (size_t)n;
  if (albedo_mix({l.x, l.y}, transfer, massHeight, debrisHeight, ld3(albedoFluvial, i), ld3(albedoDebris, i), s,
                 [&] { return ld3(albedo_bedrock, i); }, [&] { return ld3(albedo_surface, i); }, out))
    st3(albedo_surface, i, out);
}

__device__ __forceinline__ f2 ldl(const float2* __restrict__ layers, size_t i) {
  const float2 v = __ldg(layers + i);
  return {v.x, v.y};
}

__global__ void __launch_bounds__(256) k_creep(float2* __restrict__ delta, const float2* __restrict__ layers, int nx,
                                               int ny, scale3 s, float critSlope) {
  const int y = blockIdx.x * 32 + (threadIdx.x & 31);
  const int x = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= nx || y >= ny) return;
  const size_t n = (size_t)x * ny + y;
  const f2 l00 = ldl(layers, n);
  const f2 ln0 = x > 0 ? ldl(layers, n - ny) : l00;
  const f2 lp0 = x + 1 < nx ? ldl(layers, n + ny) : l00;
  const f2 l0n = y > 0 ? ldl(layers, n - 1) : l00;
  const f2 l0p = y + 1 < ny ? ldl(layers, n + 1) : l00;
  float2 d = delta[n];
  d.y += creep_cell(l00, ln0, lp0, l0n, l0p, s, critSlope);
  delta[n] = d;
}

// mass_creep as a sweep (flow.cu's shape of the work): a warp goes down a 128-column strip, a lane owns 4 cells
// (two 16-byte loads of `layers` per row, two of `delta`), the rows above and below stay in registers, the row below
// is requested one iteration ahead.  Every pair of neighbours is evaluated ONCE: with hq != hc,
// creep_pair(c, q) == -creep_pair(q, c) bit for bit (the same top, bottom and sediment enter the same expression),
// and with hq == hc both are zeros when critSlope * dx >= 0 - as is the reference's pairing of a border cell with
// itself.  So a lane computes 4 edges down and 5 edges along the row per 4 cells instead of 16 pairs (129 -> ~50
// instructions per cell), and a cell sums its four contributions in the reference's order (x+1, x-1, y+1, y-1;
// zeros of either sign do not change the sum, which starts from +0).  The identity needs finite heights: a quad whose
// window holds a NaN or an infinity is redone cell by cell by creep_exact; a negative critSlope or scale takes the
// one-thread-per-cell kernel (launcher).
constexpr int kCreepWarps = 4;
constexpr int kCreepBand = 32;

__device__ __noinline__ float creep_exact(const float2* __restrict__ layers, int x, int y, int nx, int ny, scale3 s,
                                          float critSlope) {
  const size_t n = (size_t)x * ny + y;
  const f2 l00 = ldl(layers, n);
  const f2 ln0 = x > 0 ? ldl(layers, n - ny) : l00;
  const f2 lp0 = x + 1 < nx ? ldl(layers, n + ny) : l00;
  const f2 l0n = y > 0 ? ldl(layers, n - 1) : l00;
  const f2 l0p = y + 1 < ny ? ldl(layers, n + 1) : l00;
  return creep_cell(l00, ln0, lp0, l0n, l0p, s, critSlope);
}

struct creep_raw {  // one row as loaded: cells y0 .. y0+3 as {b0, s0, b1, s1}, {b2, s2, b3, s3}; the halo cells
  float4 a, b;
  float2 l, r;
};
struct creep_row {  // heights in metres and sediment of columns y0-1 .. y0+4
  float h[6], s[6];
};

__global__ void __launch_bounds__(kCreepWarps * 32) k_creep_sweep(float2* __restrict__ delta,
                                                                  const float2* __restrict__ layers, int nx, int ny,
                                                                  int band_rows, scale3 s, float critSlope) {
  const int lane = threadIdx.x & 31;
  const int strips = (ny + 127) / 128;
  const long long wid = (long long)blockIdx.x * kCreepWarps + (threadIdx.x >> 5);
  const int x0 = (int)(wid / strips) * band_rows;
  const int y0 = (int)(wid % strips) * 128 + lane * 4;
  if (x0 >= nx || y0 >= ny) return;
  const int x1 = min(x0 + band_rows, nx);
  const bool okL = y0 > 0, okR = y0 + 4 < ny;
  const float sz = s.z, cdx = critSlope * s.x, cdy = critSlope * s.y;  // erosion.cu:676 `critSlope * dx`, rounded once

  auto fetch = [&](int x) {  // nothing may touch the values here: the loads run one iteration ahead
    const float2* p = layers + (x * ny + y0);  // nx * ny < 2^31 (launcher)
    creep_raw q;
    q.a = __ldg(reinterpret_cast<const float4*>(p));
    q.b = __ldg(reinterpret_cast<const float4*>(p + 2));
    q.l = okL ? __ldg(p - 1) : make_float2(0.0f, 0.0f);
    q.r = okR ? __ldg(p + 4) : make_float2(0.0f, 0.0f);
    return q;
  };
  auto settle = [&](const creep_raw& q) {  // erosion.cu:660-664: h = (bedrock + sediment) * sz
    creep_row r;
    r.h[0] = (q.l.x + q.l.y) * sz; r.s[0] = q.l.y;
    r.h[1] = (q.a.x + q.a.y) * sz; r.s[1] = q.a.y;
    r.h[2] = (q.a.z + q.a.w) * sz; r.s[2] = q.a.w;
    r.h[3] = (q.b.x + q.b.y) * sz; r.s[3] = q.b.y;
    r.h[4] = (q.b.z + q.b.w) * sz; r.s[4] = q.b.w;
    r.h[5] = (q.r.x + q.r.y) * sz; r.s[5] = q.r.y;
    return r;
  };
  // creep_pair (erosion_math.cuh) on heights and sediments: what cell c receives from (+) or gives to (-) neighbour q
  auto pair = [&](float sc, float sq, float hc, float hq, float cd) {
    const bool up = hq > hc;
    const float ht = up ? hq : hc;
    const float hb = up ? hc : hq;
    const float sed_top = up ? sq : sc;
    const float tmax = 0.5f * ((ht - hb) - cd);
    const float give = fmaxf(0.0f, fminf(sed_top * sz, tmax));
    return up ? give : -give;
  };

  creep_row cu = settle(fetch(x0));
  float upe[4];  // edge between rows x-1 and x as seen from row x-1; row x receives its negation
  bool finU = true;
  if (x0 > 0) {
    const creep_row ab = settle(fetch(x0 - 1));
#pragma unroll
    for (int j = 0; j < 4; ++j) upe[j] = pair(ab.s[j + 1], cu.s[j + 1], ab.h[j + 1], cu.h[j + 1], cdx);
    const float z = (((ab.h[1] + ab.h[2]) + (ab.h[3] + ab.h[4]))) * 0.0f;
    finU = z == z;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) upe[j] = 0.0f;  // the top row pairs with itself: a zero
  }
  const int xlast = min(x1, nx - 1);
  creep_raw nxt = fetch(min(x0 + 1, xlast));
#pragma unroll 2
  for (int x = x0; x < x1; ++x) {
    const creep_raw ahead = fetch(min(x + 2, xlast));
    float4* dp = reinterpret_cast<float4*>(delta + (x * ny + y0));
    float4 d0 = dp[0], d1 = dp[1];
    const creep_row dn = settle(nxt);
    const bool below = x + 1 < nx;
    float dwn[4], hz[5];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      dwn[j] = below ? pair(cu.s[j + 1], dn.s[j + 1], cu.h[j + 1], dn.h[j + 1], cdx) : 0.0f;
#pragma unroll
    for (int i = 0; i < 5; ++i) hz[i] = pair(cu.s[i], cu.s[i + 1], cu.h[i], cu.h[i + 1], cdy);
    if (!okL) hz[0] = 0.0f;  // the map ends: the cell pairs with itself, a zero
    if (!okR) hz[4] = 0.0f;
    float res[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float t = 0.0f;
      t += dwn[j];      // x+1
      t += -upe[j];     // x-1
      t += hz[j + 1];   // y+1
      t += -hz[j];      // y-1
      res[j] = 0.25f * t / sz;
    }
    // finite heights everywhere in the window?  (x * 0 is NaN for NaN and infinities)
    const float zc = ((cu.h[1] + cu.h[2]) + (cu.h[3] + cu.h[4])) * 0.0f;
    const float ze = ((okL ? cu.h[0] : 0.0f) + (okR ? cu.h[5] : 0.0f)) * 0.0f;
    const float zd = below ? ((dn.h[1] + dn.h[2]) + (dn.h[3] + dn.h[4])) * 0.0f : 0.0f;
    const bool finC = zc == zc, finD = zd == zd;
    if (!(finU && finC && finD && ze == ze)) {
      for (int j = 0; j < 4; ++j) res[j] = creep_exact(layers, x, y0 + j, nx, ny, s, critSlope);
    }
    d0.y += res[0]; d0.w += res[1]; d1.y += res[2]; d1.w += res[3];
    dp[0] = d0; dp[1] = d1;
#pragma unroll
    for (int j = 0; j < 4; ++j) upe[j] = dwn[j];
    finU = finC;
    cu = dn;
    nxt = ahead;
  }
}

__global__ void __launch_bounds__(256) k_layer_merge(float* __restrict__ height, const float2* __restrict__ layers,
                                                     long long n_cells) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_cells) return;
  const float2 l = layers[n];
  height[n] = l.x + l.y;
}

// ---- the fused step -------------------------------------------------------------
//
// Block = 32 columns x 8 rows of cells, one thread per cell; a warp covers one
// 32-cell row segment = exactly one `touched` tag.  Neighbour layers come
// through the read-only path (each line is shared by the rows above/below in
// the same block and hits L1).  ALBEDO / EMA are compile-time variants.

struct fused_args {
  int nx, ny;                    // local rows allocated, row length
  int row_lo, row_hi;            // local rows this call updates
  int row0, nx_g;                // global row of local row 0, global row count
  scale3 s;
  int with_creep;
  const float2* layers;
  float2* layers_next;
  float* height;
  const float *rainfall, *uplift;
  float *waterHeight, *mass;
  float2* momentum;
  float* debris;
  float2* debrisVel;
  const float* albedoBedrock;
  float* albedoSurface;
  // packed flux records (transport.cu fused_io): fluvial [N*R] then debris [N*R]
  float *frec, *drec;
  unsigned char* touched;            // one tag bit per cell
  float *p_frec, *p_drec;            // previous step's records (null = in-place mode)
  const unsigned char* touched_prev;
  double* mass_sum;
  float4* pre;                       // k_pre's scratch (or null): float2 gradient per cell, then float creep term per cell
  float2* pre_g;                     //   (12 of the 16 bytes per cell the caller provides: 4 bytes less to write and to read)
  float* pre_c;
  // packed particle view (transport.cu): fluvial {h, wh, vx, vy}, debris {h, vx, vy, debris} in 2 x 2-cell quads
  float4 *view_f, *view_d;
  int nyq;                           // quads per row pair
};

__device__ __forceinline__ size_t quad_of(int lx, int y, int nyq) {
  return ((size_t)(lx >> 1) * nyq + (y >> 1)) * 4 + ((lx & 1) << 1) + (y & 1);
}

// k_pre is launched with 32 KB of dynamic shared memory that it never touches.  Measured at 8192^2 (round 2): the
// particle phase it runs under takes 2.24 ms instead of 2.53 ms (any request from 16 to 40 KB does; 0.89 instead of
// 0.98 ms with 1024 particles, 1.61 instead of 1.83 ms with 16384, unchanged from 65536 up; the dense pass is
// unaffected).  The particle kernel uses no shared memory and lives on its L1 (a maximum-shared carve-out takes it from
// 2.5 to 5.7 ms); a co-resident kernel that does ask for shared memory gets fewer blocks next to it, which spreads
// k_pre's 2.5 GB of streaming traffic over the particle kernel's whole run instead of its first third.  Requests of
// 70 KB and more throttle k_pre so far that the dense pass has to wait for it (3.1 ms at 90 KB).
#ifndef SOILB200_PRE_SMEM
#define SOILB200_PRE_SMEM (32 * 1024)
#endif
constexpr int kPreSmem = SOILB200_PRE_SMEM;

// One thread that lets `ns` nanoseconds pass (strips: between the start list and k_pre on the side stream, see the
// launch below)
__global__ void k_hold(unsigned ns) {
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    __nanosleep(500);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  } while (t - t0 < ns);
}
constexpr unsigned kPreHoldNs = 20000;

// The part of the dense pass that depends on the pre-update layers only: downhill gradient and creep term.
// Launched on a side stream so that it runs WHILE the particle kernel (latency-bound, ~20 % of the issue slots
// busy) is in flight; the fused kernel then needs no neighbour loads, no gradient and no creep arithmetic.
__global__ void __launch_bounds__(256) k_pre(fused_args a, soilb200_param_t p) {
  const int y = blockIdx.x * 32 + (threadIdx.x & 31);
  const int x = a.row_lo + blockIdx.y * 8 + (threadIdx.x >> 5);
  const int ny = a.ny;
  if (x >= a.row_hi || y >= ny) return;
  const size_t n = (size_t)x * ny + y;
  const scale3 s = a.s;
  const f2 l00 = ldl(a.layers, n);
  const int gx = x + a.row0;
  const bool hxm = gx > 0, hxp = gx + 1 < a.nx_g, hym = y > 0, hyp = y + 1 < ny;
  const f2 ln0 = hxm ? ldl(a.layers, n - ny) : l00;
  const f2 lp0 = hxp ? ldl(a.layers, n + ny) : l00;
  const f2 l0n = hym ? ldl(a.layers, n - 1) : l00;
  const f2 l0p = hyp ? ldl(a.layers, n + 1) : l00;
  const float cr = a.with_creep ? creep_cell(l00, ln0, lp0, l0n, l0p, s, p.critSlopeSediment) : 0.0f;
  const f2 grad = glocal_h(l00.x + l00.y, hxm ? ln0.x + ln0.y : CUDART_NAN_F, hxp ? lp0.x + lp0.y : CUDART_NAN_F,
                           hym ? l0n.x + l0n.y : CUDART_NAN_F, hyp ? l0p.x + l0p.y : CUDART_NAN_F, s, p.exitSlope);
  a.pre_g[n] = make_float2(grad.x, grad.y);
  a.pre_c[n] = cr;
  // the flux records the previous step consumed are re-zeroed here (pure stores), off the critical path
  if (a.p_frec && ((__ldg(reinterpret_cast<const unsigned*>(a.touched_prev) + (n >> 5)) >> (n & 31)) & 1u)) {
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const int R = a.albedoSurface ? 8 : 4;
    float4* fr = reinterpret_cast<float4*>(a.p_frec + (size_t)2 * R * n);
    float4* dr = reinterpret_cast<float4*>(a.p_drec + (size_t)2 * R * n);
    fr[0] = zero4; dr[0] = zero4;
    if (R == 8) { fr[1] = zero4; dr[1] = zero4; }
  }
}

template <bool ALBEDO, bool EMA, bool PINGPONG, bool PRE, bool VIEW = false>
__global__ void __launch_bounds__(256, 5) k_fused_step(fused_args a, derived_t dv, soilb200_param_t p) {
  const int lane = threadIdx.x & 31;
  // the block covers 8 rows x 32 columns.  One warp = one row of 32 cells; in view mode one warp = 2 rows x 16 columns
  // with the lanes in quad order (lane = 4 * quad + 2 * row + column parity), so that the warp's 32 records of a
  // particle system are 512 consecutive bytes of the view - one coalesced store - while every row-major map is still
  // read and written in 64-byte runs
  const int wrp = threadIdx.x >> 5;
  const int y = blockIdx.x * 32 + (VIEW ? 16 * (wrp >> 2) + 2 * (lane >> 2) + (lane & 1) : lane);
  const int x = a.row_lo + blockIdx.y * 8 + (VIEW ? 2 * (wrp & 3) + ((lane >> 1) & 1) : wrp);
  const int ny = a.ny;
  const bool in = (x < a.row_hi) && (y < ny);
  float bed = 0.0f, sed = 0.0f;

  if (in) {
    const size_t n = (size_t)x * ny + y;
    const scale3 s = a.s;

    // the 5-point neighbourhood of the PRE-update layers
    const f2 l00 = ldl(a.layers, n);
    // map edges are GLOBAL edges: inside a strip the rows above/below are halo rows of the neighbour
    const int gx = x + a.row0;
    const bool hxm = gx > 0, hxp = gx + 1 < a.nx_g, hym = y > 0, hyp = y + 1 < ny;
    f2 ln0 = l00, lp0 = l00, l0n = l00, l0p = l00;
    float4 pre = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (PRE) {
      const float2 pg = __ldg(a.pre_g + n);  // gradient and creep term from k_pre: no neighbour loads here
      pre = make_float4(pg.x, pg.y, __ldg(a.pre_c + n), 0.0f);
    } else {
      if (hxm) ln0 = ldl(a.layers, n - ny);
      if (hxp) lp0 = ldl(a.layers, n + ny);
      if (hym) l0n = ldl(a.layers, n - 1);
      if (hyp) l0p = ldl(a.layers, n + 1);
    }
    // ---- every remaining load that depends on nothing, issued before anything is consumed ----
    constexpr int R = ALBEDO ? 8 : 4;
    const unsigned tag = (__ldg(reinterpret_cast<const unsigned*>(a.touched) + (n >> 5)) >> (n & 31)) & 1u;
    const unsigned tag_prev =
        (PINGPONG && !PRE) ? (__ldg(reinterpret_cast<const unsigned*>(a.touched_prev) + (n >> 5)) >> (n & 31)) & 1u : 0u;
    const float rain = __ldg(a.rainfall + n);
    const float upl = __ldg(a.uplift + n);
    float o_wh = 0.0f, o_ms = 0.0f, o_db = 0.0f;
    float2 o_mo = make_float2(0.0f, 0.0f), o_dv = make_float2(0.0f, 0.0f);
    // view mode: the cell's records in the packed particle view (2 x 2-cell quads) replace five of the maps
    const size_t q = VIEW ? quad_of(x, y, a.nyq) : 0;
    if (EMA && VIEW) {
      const float4 vf = a.view_f[q], vd = a.view_d[q];  // {h, wh, vx, vy}, {h, vx, vy, debris}
      o_wh = vf.y; o_ms = a.mass[n]; o_db = vd.w;
      o_mo = make_float2(vf.z, vf.w); o_dv = make_float2(vd.y, vd.z);
    } else if (EMA) {
      o_wh = a.waterHeight[n]; o_ms = a.mass[n]; o_db = a.debris[n];
      o_mo = a.momentum[n]; o_dv = a.debrisVel[n];
    }

    // ---- loads that depend on the first wave: flux records of tagged groups, bedrock colour of bare cells ----
    // flux records: only cells tagged by a particle hold non-zero data
    const bool hot = tag != 0;
    float wF = 0.0f, mF = 0.0f, dF = 0.0f;
    float2 vF = make_float2(0.0f, 0.0f), dvF = make_float2(0.0f, 0.0f);
    f3 aF = {0.0f, 0.0f, 0.0f}, aD = {0.0f, 0.0f, 0.0f};
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (hot) {
      float4* fr = reinterpret_cast<float4*>(a.frec + (size_t)2 * R * n);
      float4* dr = reinterpret_cast<float4*>(a.drec + (size_t)2 * R * n);
      // ping-pong: the records are read-only for this kernel -> non-coherent path, freely schedulable
      const float4 f0 = PINGPONG ? __ldg(fr) : fr[0];
      const float4 d0 = PINGPONG ? __ldg(dr) : dr[0];
      wF = f0.x; mF = f0.y; vF = make_float2(f0.z, f0.w);
      dF = d0.x; dvF = make_float2(d0.y, d0.z);
      if (ALBEDO) {
        const float4 f1 = PINGPONG ? __ldg(fr + 1) : fr[1];
        const float4 d1 = PINGPONG ? __ldg(dr + 1) : dr[1];
        aF = {f1.x, f1.y, f1.z}; aD = {d1.x, d1.y, d1.z};
      }
      if (!PINGPONG) {  // in-place: read-then-zero the same lines (slow on B200)
        fr[0] = zero4; dr[0] = zero4;
        if (ALBEDO) { fr[1] = zero4; dr[1] = zero4; }
      }
    }
    f3 bedrockColour = {0.0f, 0.0f, 0.0f};
    const bool bare = ALBEDO && (l00.y == 0.0f);
    if (bare) bedrockColour = {__ldg(a.albedoBedrock + 3 * n), __ldg(a.albedoBedrock + 3 * n + 1), __ldg(a.albedoBedrock + 3 * n + 2)};

    const float h = l00.x + l00.y;
    const f2 grad = PRE ? f2{pre.x, pre.y}
                        : glocal_h(h, hxm ? ln0.x + ln0.y : CUDART_NAN_F, hxp ? lp0.x + lp0.y : CUDART_NAN_F,
                                   hym ? l0n.x + l0n.y : CUDART_NAN_F, hyp ? l0p.x + l0p.y : CUDART_NAN_F, s, p.exitSlope);

    fluvial_norm fl = normalize_fluvial(grad, rain, wF, mF, {vF.x, vF.y}, dv, p);
    debris_norm db = normalize_debris(grad, dF, {dvF.x, dvF.y}, dv, p);
    if (EMA) {  // x <- (1-lrate)*x + lrate*x_new; the reference overwrites (== lrate 1)
      const float lr = p.lrate, kr = 1.0f - p.lrate;
      fl.waterHeight = kr * o_wh + lr * fl.waterHeight;
      fl.mass = kr * o_ms + lr * fl.mass;
      fl.velocity.x = kr * o_mo.x + lr * fl.velocity.x;
      fl.velocity.y = kr * o_mo.y + lr * fl.velocity.y;
      db.mass = kr * o_db + lr * db.mass;
      db.velocity.x = kr * o_dv.x + lr * db.velocity.x;
      db.velocity.y = kr * o_dv.y + lr * db.velocity.y;
    }

    f2 delta = {0.0f, 0.0f};
    const float transfer = transfer_cell(grad, l00, delta, upl, fl.mass, fl.velocity, db.mass, s, dv, p);

    f3 newSurface = {0.0f, 0.0f, 0.0f};
    bool writeSurface = false;
    if (ALBEDO) {
      // colours the normalise kernels would have written into the flux maps (erosion.cu:181-185,387-391);
      // untouched cells fall back to albedoSource == albedo_surface, which is only needed if a mix happens
      const bool needF = !(mF > 0.0f && len2(aF.x, aF.y) > 0.0f);
      const bool needD = !(dF > 0.0f && len2(aD.x, aD.y) > 0.0f);
      const float eps = 1E-12f;
      const bool mixes = (l00.y != 0.0f) && ((fl.mass + db.mass) > 0.0f) && (transfer > eps);
      f3 surf = {0.0f, 0.0f, 0.0f};
      if (mixes) surf = ld3(a.albedoSurface, n);
      const f3 cF = needF ? surf : f3{aF.x / mF, aF.y / mF, aF.z / mF};
      const f3 cD = needD ? surf : f3{aD.x / dF, aD.y / dF, aD.z / dF};
      writeSurface = albedo_mix(l00, transfer, fl.mass, db.mass, cF, cD, s, [&] { return bedrockColour; },
                                [&] { return surf; }, newSurface);
    }

    if (PRE) {
      if (a.with_creep) delta.y += pre.z;
    } else if (a.with_creep) {
      delta.y += creep_cell(l00, ln0, lp0, l0n, l0p, s, p.critSlopeSediment);
    }

    bed = l00.x + delta.x;
    sed = l00.y + delta.y;

    // ---- all stores ----
    st_stream(a.mass + n, fl.mass);
    a.layers_next[n] = make_float2(bed, sed);
    if (VIEW) {  // what the particles of the next step gather, one 16-byte record per particle system
      a.view_f[q] = make_float4(bed + sed, fl.waterHeight, fl.velocity.x, fl.velocity.y);
      a.view_d[q] = make_float4(bed + sed, db.velocity.x, db.velocity.y, db.mass);
    } else {
      st_stream(a.waterHeight + n, fl.waterHeight);
      a.momentum[n] = make_float2(fl.velocity.x, fl.velocity.y);
      st_stream(a.debris + n, db.mass);
      a.debrisVel[n] = make_float2(db.velocity.x, db.velocity.y);
      a.height[n] = bed + sed;  // re-read by the particles of the next step
    }
    if (ALBEDO && writeSurface) st3(a.albedoSurface, n, newSurface);
    if (PINGPONG && tag_prev != 0) {  // pure stores into the records the previous step consumed
      float4* fr = reinterpret_cast<float4*>(a.p_frec + (size_t)2 * R * n);
      float4* dr = reinterpret_cast<float4*>(a.p_drec + (size_t)2 * R * n);
      fr[0] = zero4; dr[0] = zero4;
      if (ALBEDO) { fr[1] = zero4; dr[1] = zero4; }
    }
  }

  if (a.mass_sum) {  // warp-shuffle reduction, one double atomic pair per block
    double b = in ? (double)bed : 0.0, sd = in ? (double)sed : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      b += __shfl_xor_sync(0xffffffffu, b, o);
      sd += __shfl_xor_sync(0xffffffffu, sd, o);
    }
    __shared__ double sb_[8], ss_[8];
    if (lane == 0) { sb_[threadIdx.x >> 5] = b; ss_[threadIdx.x >> 5] = sd; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double tb = 0.0, ts = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) { tb += sb_[w]; ts += ss_[w]; }
      atomicAdd(a.mass_sum, tb);
      atomicAdd(a.mass_sum + 1, ts);
    }
  }
}

// maps <-> view, one thread per cell over all allocated rows (halo rows included)
__global__ void __launch_bounds__(256) k_view_import(const float2* __restrict__ layers, const float* __restrict__ wh,
                                                     const float2* __restrict__ mom, const float2* __restrict__ dmom,
                                                     const float* __restrict__ debris, float4* __restrict__ view_f,
                                                     float4* __restrict__ view_d, int nx, int ny, int nyq) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  const float2 l = layers[n];
  const float h = l.x + l.y;
  const float2 m = mom[n], d = dmom[n];
  const size_t q = quad_of(x, y, nyq);
  view_f[q] = make_float4(h, wh[n], m.x, m.y);
  view_d[q] = make_float4(h, d.x, d.y, debris[n]);
}

__global__ void __launch_bounds__(256) k_view_export(const float4* __restrict__ view_f, const float4* __restrict__ view_d,
                                                     float* __restrict__ height, float* __restrict__ wh,
                                                     float2* __restrict__ mom, float2* __restrict__ dmom,
                                                     float* __restrict__ debris, int nx, int ny, int nyq) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  const size_t q = quad_of(x, y, nyq);
  const float4 f = view_f[q], d = view_d[q];
  height[n] = f.x;
  wh[n] = f.y;
  mom[n] = make_float2(f.z, f.w);
  dmom[n] = make_float2(d.y, d.z);
  debris[n] = d.w;
}

// self-test of cdiv (common.cuh): every binary32 numerator through the launch-constant division against the IEEE one
__global__ void __launch_bounds__(256) k_selftest_div_const(float d, float r, int kind, unsigned long long* mismatches) {
  unsigned long long bad = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32);
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const float x = __uint_as_float((unsigned)i);
    const unsigned a = __float_as_uint(cdiv(x, d, r, kind)), b = __float_as_uint(x / d);
    if (a != b && x == x) ++bad;   // NaN numerators: both NaN
  }
  if (bad) atomicAdd(mismatches, bad);
}

__device__ unsigned long long d_hot_segments = 0ull;  // profiling only: tags found set, summed over steps

// tags are cleared after the dense pass has consumed them
__global__ void __launch_bounds__(256) k_clear(unsigned* __restrict__ p, long long words) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < words) p[i] = 0u;
}

// profiling only: number of set tags (a pure read pass; reading and clearing in one kernel would be the slow
// read-then-zero pattern)
__global__ void __launch_bounds__(256) k_count_tags(const unsigned* __restrict__ p, long long words) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int c = 0;
  if (i < words) {
    const unsigned w = __ldg(p + i);
    c = __popc(w);
  }
  c = __reduce_add_sync(0xffffffffu, c);
  __shared__ int part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[w];
    if (t) atomicAdd(&d_hot_segments, (unsigned long long)t);
  }
}

// optional per-phase timing (bench.py's roofline line): CUDA events recorded on the caller's
// stream around the particle kernels and around the dense kernel of every step
struct profiler {
  bool on = false;
  std::vector<cudaEvent_t> ev;  // 3 per step: start, after particles, after dense
  size_t used = 0;
  cudaEvent_t next() {
    if (used == ev.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      ev.push_back(e);
    }
    return ev[used++];
  }
} g_prof;

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

int soilb200_transport_fluvial(const float* layers, const float* rainfall, float* discharge, float* discharge_track,
                               float* mass, float* mass_track, float* momentum, float* momentum_track,
                               const float* albedo_bedrock, float* albedo_transport, const float* albedo_surface,
                               void* rng, int64_t n_particles, int nx, int ny, const float scale[3],
                               const soilb200_param_t* param, void* stream) {
  (void)albedo_bedrock;  // accepted and unused, as in the reference (App. B-6)
  if (!layers || !rainfall || !discharge || !discharge_track || !mass || !mass_track || !momentum || !momentum_track ||
      !rng || !scale || !param || nx <= 0 || ny <= 0 || n_particles < 0)
    return SOILB200_E_ARG;
  if ((albedo_transport == nullptr) != (albedo_surface == nullptr)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  const scale3 sc{scale[0], scale[1], scale[2]};
  int rc = launch_transport_fluvial(discharge_track, mass_track, momentum_track, albedo_transport, rng, n_particles,
                                    layers, rainfall, discharge, momentum, albedo_surface, nx, ny, sc, *param, s);
  if (rc) return rc;
  const long long N = (long long)nx * ny;
  k_normalize_fluvial<<<blocks_for(N, 256), 256, 0, s>>>(discharge_track, mass_track, (const float2*)momentum_track,
                                                         albedo_transport, (const float2*)layers, rainfall, discharge,
                                                         mass, (float2*)momentum, albedo_surface, nx, ny, sc, derive(*param, sc), *param);
  count_launch();
  return check_launch();
}

int soilb200_transport_debris(const float* layers, float* velocity, float* velocity_track, float* mass,
                              float* mass_track, const float* albedo_bedrock, float* albedo_transport,
                              const float* albedo_surface, void* rng, int64_t n_particles, int nx, int ny,
                              const float scale[3], const soilb200_param_t* param, void* stream) {
  (void)albedo_bedrock;
  if (!layers || !velocity || !velocity_track || !mass || !mass_track || !rng || !scale || !param || nx <= 0 ||
      ny <= 0 || n_particles < 0)
    return SOILB200_E_ARG;
  if ((albedo_transport == nullptr) != (albedo_surface == nullptr)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  const scale3 sc{scale[0], scale[1], scale[2]};
  int rc = launch_transport_debris(mass_track, velocity_track, albedo_transport, rng, n_particles, layers, velocity,
                                   albedo_surface, nx, ny, sc, *param, s);
  if (rc) return rc;
  const long long N = (long long)nx * ny;
  k_normalize_debris<<<blocks_for(N, 256), 256, 0, s>>>(mass_track, (const float2*)velocity_track, albedo_transport,
                                                        (const float2*)layers, mass, (float2*)velocity, albedo_surface,
                                                        nx, ny, sc, derive(*param, sc), *param);
  count_launch();
  return check_launch();
}

int soilb200_mass_transfer(float* delta, const float* layers, const float* uplift, const float* discharge,
                           const float* mass, const float* momentum, const float* debris,
                           const float* momentum_debris, const float* albedo_bedrock,
                           const float* albedo_transport_fluvial, const float* albedo_transport_debris,
                           float* albedo_surface, int nx, int ny, const float scale[3],
                           const soilb200_param_t* param, void* stream) {
  (void)discharge; (void)momentum_debris;  // accepted and unused, as in the reference (App. B-6)
  if (!delta || !layers || !uplift || !mass || !momentum || !debris || !scale || !param || nx <= 0 || ny <= 0)
    return SOILB200_E_ARG;
  const int na = (albedo_bedrock != nullptr) + (albedo_transport_fluvial != nullptr) +
                 (albedo_transport_debris != nullptr) + (albedo_surface != nullptr);
  if (na != 0 && na != 4) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  const long long N = (long long)nx * ny;
  k_transfer<<<blocks_for(N, 256), 256, 0, (cudaStream_t)stream>>>(
      (float2*)delta, (const float2*)layers, uplift, mass, (const float2*)momentum, debris, albedo_bedrock,
      albedo_transport_fluvial, albedo_transport_debris, albedo_surface, nx, ny, scale3{scale[0], scale[1], scale[2]},
      derive(*param, scale3{scale[0], scale[1], scale[2]}), *param);
  count_launch();
  return check_launch();
}

int soilb200_mass_creep(float* delta, const float* layers, int nx, int ny, const float scale[3],
                        const soilb200_param_t* param, void* stream) {
  if (!delta || !layers || !scale || !param || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  const float crit = param->critSlopeSediment;
  const bool aligned = ((reinterpret_cast<uintptr_t>(delta) | reinterpret_cast<uintptr_t>(layers)) & 15) == 0;
  // the sweep evaluates every pair of neighbours once; that identity needs critSlope * dx >= 0 (k_creep_sweep)
  if (ny % 4 == 0 && aligned && (long long)nx * ny <= 0x7fffffffLL && crit >= 0.0f && scale[0] > 0.0f && scale[1] > 0.0f) {
    const long long strips = (ny + 127) / 128;
    long long rows = (long long)nx * strips / ((long long)sm_count() * 48);  // few rows on small maps: fill the GPU
    rows = rows < 4 ? 4 : (rows > kCreepBand ? kCreepBand : rows);
    const long long warps = strips * ((nx + rows - 1) / rows);
    k_creep_sweep<<<blocks_for(warps, kCreepWarps), kCreepWarps * 32, 0, (cudaStream_t)stream>>>(
        (float2*)delta, (const float2*)layers, nx, ny, (int)rows, scale3{scale[0], scale[1], scale[2]}, crit);
  } else {
    const dim3 grid((ny + 31) / 32, (nx + 7) / 8);
    k_creep<<<grid, 256, 0, (cudaStream_t)stream>>>((float2*)delta, (const float2*)layers, nx, ny,
                                                    scale3{scale[0], scale[1], scale[2]}, crit);
  }
  count_launch();
  return check_launch();
}

int soilb200_layer_merge(float* height, const float* layers, int nx, int ny, void* stream) {
  if (!height || !layers || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  const long long N = (long long)nx * ny;
  k_layer_merge<<<blocks_for(N, 256), 256, 0, (cudaStream_t)stream>>>(height, (const float2*)layers, N);
  count_launch();
  return check_launch();
}

size_t soilb200_erode_ws(int nx, int ny) {
  if (nx <= 0 || ny <= 0) return 0;
  const long long bytes = ((long long)nx * ny + 7) / 8;  // one tag bit per cell
  return (size_t)((bytes + 255) / 256 * 256);
}

size_t soilb200_erode_spawn_bytes(int64_t n_particles) {
  return n_particles > 0 ? 256 + 2 * (size_t)n_particles * sizeof(float2) : 0;
}

size_t soilb200_erode_pre_bytes(int nx, int ny) {
  if (nx <= 0 || ny <= 0) return 0;
  return (size_t)nx * ny * sizeof(float4);
}

size_t soilb200_erode_view_bytes(int nx, int ny) {
  if (nx <= 0 || ny <= 0) return 0;
  // two views (fluvial, debris) of 16 bytes per cell over the map rounded up to whole 2 x 2 quads
  return (size_t)2 * ((nx + 1) / 2) * ((ny + 1) / 2) * 4 * sizeof(float4);
}

size_t soilb200_erode_flux_bytes(int nx, int ny, int albedo) {
  if (nx <= 0 || ny <= 0) return 0;
  return (size_t)nx * ny * 2 * (albedo ? 8 : 4) * sizeof(float);
}

extern "C++" {
namespace {

// side stream on which k_pre overlaps the particle kernel; created on first use, one per device.  The fork / join
// state is kept per model (keyed by its `pre` scratch, which no two models share): several strips on one GPU, or
// several models, each join their own k_pre.
struct side_slot {
  const void* pre = nullptr;
  cudaEvent_t join = nullptr;
  bool pending = false;  // a k_pre has been launched whose join the next dense pass must wait for
};
struct side_t {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr;
  side_slot slots[32];
  side_slot* slot(const void* pre) {
    side_slot* free_slot = nullptr;
    for (side_slot& sl : slots) {
      if (sl.pre == pre) return &sl;
      if (!sl.pre && !free_slot) free_slot = &sl;
    }
    if (!free_slot) {  // all in use: recycle one that has nothing in flight
      for (side_slot& sl : slots)
        if (!sl.pending) { free_slot = &sl; break; }
      if (!free_slot) return nullptr;
    }
    if (!free_slot->join && cudaEventCreateWithFlags(&free_slot->join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    free_slot->pre = pre;
    free_slot->pending = false;
    return free_slot;
  }
} g_side[64];

inline side_t* side_for_current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  side_t& sd = g_side[dev];
  if (!sd.stream) {
    if (cudaStreamCreateWithFlags(&sd.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&sd.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
  }
  return &sd;
}

// one instantiation per combination of colour channels, learning rate, ping-pong records, pre scratch use and view
template <bool A, bool E>
inline void launch_fused(bool pingpong, bool pre, bool view, dim3 grid, cudaStream_t s, const fused_args& a,
                         const derived_t& dv, const soilb200_param_t& p) {
  if (!pingpong) {
    if (pre) k_fused_step<A, E, false, true, false><<<grid, 256, 0, s>>>(a, dv, p);
    else k_fused_step<A, E, false, false, false><<<grid, 256, 0, s>>>(a, dv, p);
  } else if (view) {
    if (pre) k_fused_step<A, E, true, true, true><<<grid, 256, 0, s>>>(a, dv, p);
    else k_fused_step<A, E, true, false, true><<<grid, 256, 0, s>>>(a, dv, p);
  } else {
    if (pre) k_fused_step<A, E, true, true, false><<<grid, 256, 0, s>>>(a, dv, p);
    else k_fused_step<A, E, true, false, false><<<grid, 256, 0, s>>>(a, dv, p);
  }
}

// strip description of a model: all-zero strip fields mean "the arrays are the whole map"
inline int strip_of(const soilb200_model_t* m, strip_host& sh, int& row_lo, int& row_hi) {
  const bool whole = m->global_nx == 0;
  sh.nx_g = whole ? m->nx : m->global_nx;
  sh.row0 = whole ? 0 : m->global_row0;
  sh.own_lo = whole ? 0 : m->own_row_lo;
  sh.own_hi = whole ? m->nx : m->own_row_hi;
  sh.outbox_lo = (float*)m->outbox_lo; sh.outbox_hi = (float*)m->outbox_hi;
  sh.counts = m->outbox_counts; sh.capacity = m->outbox_capacity;
  sh.peers = m->peers; sh.n_peers = m->n_peers; sh.peer_self = m->peer_self;
  if (m->peers && (m->n_peers <= 0 || m->peer_self < 0 || m->peer_self >= m->n_peers)) return SOILB200_E_ARG;
  row_lo = sh.own_lo - sh.row0; row_hi = sh.own_hi - sh.row0;
  if (row_lo < 0 || row_hi > m->nx || row_lo >= row_hi || sh.row0 < 0 || sh.own_hi > sh.nx_g) return SOILB200_E_ARG;
  // a halo row must be allocated wherever a neighbouring strip exists
  if (sh.own_lo > 0 && row_lo < 1) return SOILB200_E_ARG;
  if (sh.own_hi < sh.nx_g && row_hi > m->nx - 1) return SOILB200_E_ARG;
  const bool has_neighbour = sh.own_lo > 0 || sh.own_hi < sh.nx_g;
  if (has_neighbour && !sh.peers && (!sh.outbox_lo || !sh.outbox_hi || !sh.counts || sh.capacity <= 0)) return SOILB200_E_ARG;
  if ((reinterpret_cast<uintptr_t>(sh.outbox_lo) & 15) || (reinterpret_cast<uintptr_t>(sh.outbox_hi) & 15)) return SOILB200_E_ARG;
  return SOILB200_OK;
}

// the fused step divides by its launch constants with cdiv (common.cuh); SOILB200_PLAIN_DIV=1 keeps the plain
// divisions (for A/B measurements)
inline scale3 fused_scale(const float scale[3]) {
  static const bool plain = [] { const char* e = getenv("SOILB200_PLAIN_DIV"); return e && e[0] == '1'; }();
  return make_scale(scale[0], scale[1], scale[2], !plain);
}

inline int check_model(const soilb200_model_t* m, const soilb200_param_t* param) {
  if (!m || !param) return SOILB200_E_ARG;
  if (m->nx <= 0 || m->ny <= 0 || m->n_particles < 0) return SOILB200_E_ARG;
  if (!m->layers || !m->layers_next || !m->height || !m->rainfall || !m->uplift || !m->discharge || !m->mass ||
      !m->momentum || !m->debris || !m->debris_momentum || !m->rng_fluvial || !m->rng_debris || !m->flux ||
      !m->touched)
    return SOILB200_E_ARG;
  const int na = (m->albedo_bedrock != nullptr) + (m->albedo_surface != nullptr);
  if (na != 0 && na != 2) return SOILB200_E_ARG;
  if ((m->flux_prev != nullptr) != (m->touched_prev != nullptr)) return SOILB200_E_ARG;
  if ((long long)m->nx * m->ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  if ((long long)(m->global_nx ? m->global_nx : m->nx) * m->ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  if ((reinterpret_cast<uintptr_t>(m->flux) & 15) || (reinterpret_cast<uintptr_t>(m->flux_prev) & 15)) return SOILB200_E_ARG;
  return SOILB200_OK;
}

}  // namespace
}  // extern "C++"

int soilb200_erode_step(const soilb200_model_t* m, const soilb200_param_t* param, void* stream) {
  int rc = check_model(m, param);
  if (rc) return rc;
  strip_host sh;
  int row_lo, row_hi;
  rc = strip_of(m, sh, row_lo, row_hi);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const scale3 sc = fused_scale(m->scale);
  unsigned char* touched = (unsigned char*)m->touched;
  const bool albedo = m->albedo_surface != nullptr;
  const bool pingpong = m->flux_prev != nullptr;
  const int R = albedo ? 8 : 4;
  const int phases = m->phases == 0 ? (SOILB200_PHASE_PARTICLES | SOILB200_PHASE_DENSE) : m->phases;

  fused_args a;
  a.nx = m->nx; a.ny = m->ny; a.s = sc; a.with_creep = m->with_creep;
  a.row_lo = row_lo; a.row_hi = row_hi; a.row0 = sh.row0; a.nx_g = sh.nx_g;
  a.layers = (const float2*)m->layers; a.layers_next = (float2*)m->layers_next; a.height = m->height;
  a.rainfall = m->rainfall; a.uplift = m->uplift;
  a.waterHeight = m->discharge; a.mass = m->mass; a.momentum = (float2*)m->momentum;
  a.debris = m->debris; a.debrisVel = (float2*)m->debris_momentum;
  a.albedoBedrock = m->albedo_bedrock; a.albedoSurface = m->albedo_surface;
  // a cell's fluvial and debris records are adjacent (2R floats per cell): the dense pass reads both, and a
  // scattered 32-byte read costs a 64-byte DRAM access anyway
  a.frec = (float*)m->flux; a.drec = (float*)m->flux + R;
  a.p_frec = (float*)m->flux_prev; a.p_drec = pingpong ? (float*)m->flux_prev + R : nullptr;
  a.touched = touched; a.touched_prev = (const unsigned char*)m->touched_prev;
  a.mass_sum = m->mass_sum;
  a.pre = (float4*)m->pre;
  a.pre_g = (float2*)m->pre;
  a.pre_c = m->pre ? (float*)m->pre + (size_t)2 * m->nx * m->ny : nullptr;
  a.nyq = (m->ny + 1) / 2;
  a.view_f = (float4*)m->view;
  a.view_d = m->view ? (float4*)m->view + (size_t)((m->nx + 1) / 2) * a.nyq * 4 : nullptr;
  if (m->view) {  // view mode is built on the ping-pong flux records
    if (!pingpong || (reinterpret_cast<uintptr_t>(m->view) & 15)) return SOILB200_E_ARG;
  }
  const dim3 grid((m->ny + 31) / 32, (row_hi - row_lo + 7) / 8);   // k_pre: all owned rows
  side_t* side = m->pre ? side_for_current_device() : nullptr;
  side_slot* slot = side ? side->slot(m->pre) : nullptr;
  if (m->pre && (!side || !slot || (reinterpret_cast<uintptr_t>(m->pre) & 15))) return SOILB200_E_ARG;

  if (phases & SOILB200_PHASE_PARTICLES) {
    if (g_prof.on && g_prof.used % 3 == 0) cudaEventRecord(g_prof.next(), s);
    // the start positions first: k_pre, forked below, must not flood the GPU before the particle kernel is in
    rc = launch_spawn(m->rng_fluvial, m->rng_debris, m->n_particles, m->ny, sh, m->spawn, s);
    if (rc) return rc;
    if (side) cudaEventRecord(side->fork, s);  // k_pre depends on everything before the particles, not on them
    rc = launch_transport_both(m->height, (float*)m->flux, R, touched, m->rng_fluvial, m->rng_debris, m->n_particles,
                               m->rainfall, m->discharge, m->momentum, m->debris_momentum, m->albedo_surface, m->nx,
                               m->ny, sc, *param, sh, m->spawn, m->view, s);
    if (rc) return rc;
    if (side) {  // fork: gradient + creep of the pre-update layers on the side stream, under the particle kernel
      cudaStreamWaitEvent(side->stream, side->fork, 0);
      // A strip's particle kernel follows k_spawn on its stream, and k_pre waits for the same k_spawn: both become
      // runnable at the same instant, and when k_pre's blocks reach the SMs first the particle kernel starts with the
      // slots they leave (2.38 instead of 2.15 ms as rank 0 of 8; a whole map's particle kernel is launched first and
      // has no such race: no difference measured).  k_pre has a millisecond of slack: it starts 20 us later.
      if (sh.own_lo > 0 || sh.own_hi < sh.nx_g) {
        k_hold<<<1, 1, 0, side->stream>>>(kPreHoldNs);
        count_launch();
      }
      k_pre<<<grid, 256, kPreSmem, side->stream>>>(a, *param);
      cudaEventRecord(slot->join, side->stream);
      slot->pending = true;
      count_launch();
    }
  }
  if (!(phases & SOILB200_PHASE_DENSE)) return check_launch();

  const bool cont = (phases & SOILB200_PHASE_DENSE_CONT) != 0, more = (phases & SOILB200_PHASE_DENSE_MORE) != 0;
  if (m->mass_sum && !cont) {
    rc = (int)cudaMemsetAsync(m->mass_sum, 0, 2 * sizeof(double), s);
    if (rc) return rc;
  }
  bool use_pre = false;
  if (side) {
    if (cont) {
      // an earlier dense call of this step has joined (or computed) the record
    } else if (!slot->pending) {  // dense-only call without a preceding particle phase: compute it now, in order
      k_pre<<<grid, 256, kPreSmem, s>>>(a, *param);
      count_launch();
    } else {
      cudaStreamWaitEvent(s, slot->join, 0);  // join
      slot->pending = false;
    }
    use_pre = true;
  }
  // a dense pass in several calls: this one covers the owned rows inside [dense_row_lo, dense_row_hi) (global)
  dim3 dgrid = grid;
  if (m->dense_row_lo < m->dense_row_hi) {
    const int lo = max(m->dense_row_lo - sh.row0, row_lo), hi = min(m->dense_row_hi - sh.row0, row_hi);
    if (lo >= hi) return SOILB200_E_ARG;
    a.row_lo = lo; a.row_hi = hi;
    dgrid = dim3((m->ny + 31) / 32, (hi - lo + 7) / 8);
  }

  const bool ema = param->lrate != 1.0f;
  // a step split over two calls (strips: particles, barrier, dense) is timed the same way: the first interval
  // then also holds whatever the caller enqueued between the two phases
  // (a dense pass in several calls is timed from its first call to its last)
  if (g_prof.on && !cont && g_prof.used % 3 == 1) cudaEventRecord(g_prof.next(), s);
  const derived_t dv = derive(*param, sc);
  const bool pre_kind = use_pre;
  if (albedo) {
    if (ema) launch_fused<true, true>(pingpong, pre_kind, m->view != nullptr, dgrid, s, a, dv, *param);
    else launch_fused<true, false>(pingpong, pre_kind, m->view != nullptr, dgrid, s, a, dv, *param);
  } else {
    if (ema) launch_fused<false, true>(pingpong, pre_kind, m->view != nullptr, dgrid, s, a, dv, *param);
    else launch_fused<false, false>(pingpong, pre_kind, m->view != nullptr, dgrid, s, a, dv, *param);
  }
  if (more) return check_launch();
  if (g_prof.on && g_prof.used % 3 == 2) cudaEventRecord(g_prof.next(), s);  // the step's last dense call clears the spent tags
  const long long words = (long long)(soilb200_erode_ws(m->nx, m->ny) / 4);
  // in-place mode consumed this step's tags; ping-pong mode consumed the previous step's
  unsigned* spent = (unsigned*)(pingpong ? m->touched_prev : m->touched);
  if (g_prof.on) k_count_tags<<<blocks_for(words, 256), 256, 0, s>>>(spent, words);
  k_clear<<<blocks_for(words, 256), 256, 0, s>>>(spent, words);
  count_launch(2);
  return check_launch();
}

int soilb200_erode_resume(const soilb200_model_t* m, const soilb200_param_t* param, const void* inbox, int n_in,
                          void* stream) {
  int rc = check_model(m, param);
  if (rc) return rc;
  if (n_in < 0 || (n_in > 0 && !inbox) || (reinterpret_cast<uintptr_t>(inbox) & 15)) return SOILB200_E_ARG;
  strip_host sh;
  int row_lo, row_hi;
  rc = strip_of(m, sh, row_lo, row_hi);
  if (rc) return rc;
  const scale3 sc = fused_scale(m->scale);
  return launch_transport_resume(m->height, (float*)m->flux, m->albedo_surface ? 8 : 4, (unsigned char*)m->touched,
                                 (const float*)inbox, n_in, m->n_particles, m->discharge, m->momentum,
                                 m->debris_momentum, m->nx, m->ny, sc, *param, sh, m->view, (cudaStream_t)stream);
}

int soilb200_erode_import(const soilb200_model_t* m, void* stream) {
  if (!m || !m->view || !m->layers || !m->discharge || !m->momentum || !m->debris || !m->debris_momentum || m->nx <= 0 ||
      m->ny <= 0)
    return SOILB200_E_ARG;
  if ((long long)m->nx * m->ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  const int nyq = (m->ny + 1) / 2;
  float4* vf = (float4*)m->view;
  float4* vd = vf + (size_t)((m->nx + 1) / 2) * nyq * 4;
  const long long N = (long long)m->nx * m->ny;
  k_view_import<<<blocks_for(N, 256), 256, 0, (cudaStream_t)stream>>>(
      (const float2*)m->layers, m->discharge, (const float2*)m->momentum, (const float2*)m->debris_momentum, m->debris, vf,
      vd, m->nx, m->ny, nyq);
  count_launch();
  return check_launch();
}

int soilb200_erode_export(const soilb200_model_t* m, void* stream) {
  if (!m || !m->view || !m->height || !m->discharge || !m->momentum || !m->debris || !m->debris_momentum || m->nx <= 0 ||
      m->ny <= 0)
    return SOILB200_E_ARG;
  if ((long long)m->nx * m->ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  const int nyq = (m->ny + 1) / 2;
  const float4* vf = (const float4*)m->view;
  const float4* vd = vf + (size_t)((m->nx + 1) / 2) * nyq * 4;
  const long long N = (long long)m->nx * m->ny;
  k_view_export<<<blocks_for(N, 256), 256, 0, (cudaStream_t)stream>>>(vf, vd, m->height, m->discharge, (float2*)m->momentum,
                                                                     (float2*)m->debris_momentum, m->debris, m->nx, m->ny,
                                                                     nyq);
  count_launch();
  return check_launch();
}

int soilb200_enable_peer_access(int peer_device) {
  int cur = -1;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return (int)e;
  if (peer_device == cur) return SOILB200_OK;
  int can = 0;
  e = cudaDeviceCanAccessPeer(&can, cur, peer_device);
  if (e != cudaSuccess) return (int)e;
  if (!can) return SOILB200_E_LIMIT;
  e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return SOILB200_OK; }
  return (int)e;
}

int soilb200_ipc_export(const void* ptr, unsigned char handle[64], uint64_t* offset) {
  if (!ptr || !handle || !offset) return SOILB200_E_ARG;
  // the handle names the whole allocation `ptr` lives in (torch sub-allocates): find its base
  typedef CUresult (*range_fn)(CUdeviceptr*, size_t*, CUdeviceptr);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess) return (int)e;
  if (!fn || q != cudaDriverEntryPointSuccess) return SOILB200_E_LIMIT;
  CUdeviceptr base = 0;
  size_t size = 0;
  if (((range_fn)fn)(&base, &size, (CUdeviceptr)ptr) != CUDA_SUCCESS) return SOILB200_E_ARG;
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, (void*)base);
  if (e != cudaSuccess) return (int)e;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle");
  std::memcpy(handle, &h, 64);
  *offset = (uint64_t)((CUdeviceptr)ptr - base);
  return SOILB200_OK;
}

int soilb200_ipc_open(const unsigned char handle[64], void** base) {
  if (!handle || !base) return SOILB200_E_ARG;
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, 64);
  return (int)cudaIpcOpenMemHandle(base, h, cudaIpcMemLazyEnablePeerAccess);
}

int soilb200_ipc_close(void* base) {
  if (!base) return SOILB200_E_ARG;
  return (int)cudaIpcCloseMemHandle(base);
}

int soilb200_selftest_div_const(float divisor, uint64_t* mismatches_device, int32_t* kind, void* stream) {
  if (!mismatches_device) return SOILB200_E_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(mismatches_device, 0, sizeof(uint64_t), s);
  if (e != cudaSuccess) return (int)e;
  float r = 0.0f;
  const int k = divisor_kind(divisor, &r);
  if (kind) *kind = k;
  k_selftest_div_const<<<sm_count() * 16, 256, 0, s>>>(divisor, r, k, (unsigned long long*)mismatches_device);
  count_launch();
  return check_launch();
}

int soilb200_profile_enable(int on) {
  g_prof.on = on != 0;
  g_prof.used = 0;
  const unsigned long long zero = 0ull;
  return (int)cudaMemcpyToSymbol(d_hot_segments, &zero, sizeof(zero));
}

int soilb200_profile_read(double* particle_ms, double* dense_ms, int64_t* steps, uint64_t* hot_segments) {
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return (int)e;
  double pm = 0.0, dm = 0.0;
  const size_t n = g_prof.used / 3;
  for (size_t i = 0; i < n; ++i) {
    float a = 0.0f, b = 0.0f;
    cudaEventElapsedTime(&a, g_prof.ev[3 * i], g_prof.ev[3 * i + 1]);
    cudaEventElapsedTime(&b, g_prof.ev[3 * i + 1], g_prof.ev[3 * i + 2]);
    pm += a;
    dm += b;
  }
  unsigned long long hot = 0ull;
  e = cudaMemcpyFromSymbol(&hot, d_hot_segments, sizeof(hot));
  if (particle_ms) *particle_ms = pm;
  if (dense_ms) *dense_ms = dm;
  if (steps) *steps = (int64_t)n;
  if (hot_segments) *hot_segments = hot;
  g_prof.used = 0;
  return (int)e;
}

}  // extern "C"
