// flow.cu - D4/D8 steepest receiver, direction code, Gibbs-random receiver and
// the normal map: radius-1 (radius-2 for normal) stencils over one float map.
//
// Reference: source/soillib/model/graph/graph.cu:27-264, op/normal.hpp:18-39.
//
// B200 shape of the work.  Each warp sweeps a 128-column strip down a band of
// rows.  A lane owns 4 consecutive cells (one 16-byte load per row); the left
// and right halo cells are the neighbouring lanes' lines (L1 hits; warp
// shuffles in the Gibbs sweep), the rows above and below stay in registers
// from the previous / for the next iteration, so every height is fetched from
// L2/HBM once per band (+2 halo rows per band; +4 for normal).  Outputs leave
// as 16-byte stores.  Algorithmic traffic 8 B/cell (16 for normal).  steepest /
// direction additionally share every edge's slope between its two cells
// (k_steepest).  A scalar one-thread-per-cell kernel covers maps whose row
// length is not a multiple of 4 or whose base is not 16-byte aligned.
#include "common.cuh"
#include <cstdlib>
#include <cuda.h>
#include <curand_kernel.h>

namespace sb {

unsigned long long g_launches = 0;

namespace {

#ifndef SOILB200_BAND_ROWS
#define SOILB200_BAND_ROWS 32
#endif
#ifndef SOILB200_SWEEP_WARPS
#define SOILB200_SWEEP_WARPS 4
#endif
constexpr int kBandRows = SOILB200_BAND_ROWS;         // rows swept by one warp
constexpr int kWarpsPerBlock = SOILB200_SWEEP_WARPS;  // 128 threads
constexpr float kSqrt2 = 1.41421354f;  // sqrtf(2.0f) == __length(ivec2(1,1)), graph.cu:23-25,56

struct row6 {  // one row of the register window: left halo, 4 owned cells, right halo
  float l, a, b, c, d, r;
};

__device__ __forceinline__ row6 load_row(const float* __restrict__ h, int x, int y0, int nx, int ny, int lane,
                                         bool active) {
  row6 w;
  w.l = w.a = w.b = w.c = w.d = w.r = CUDART_NAN_F;
  const bool rowok = (x >= 0) && (x < nx);
  if (rowok && active) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(h + (size_t)x * ny + y0));
    w.a = v.x; w.b = v.y; w.c = v.z; w.d = v.w;
  }
  const float fromL = __shfl_up_sync(0xffffffffu, w.d, 1);
  const float fromR = __shfl_down_sync(0xffffffffu, w.a, 1);
  w.l = fromL;
  w.r = fromR;
  if (rowok && active) {
    if (lane == 0) w.l = (y0 > 0) ? __ldg(h + (size_t)x * ny + y0 - 1) : CUDART_NAN_F;
    if (lane == 31 || y0 + 4 >= ny) w.r = (y0 + 4 < ny) ? __ldg(h + (size_t)x * ny + y0 + 4) : CUDART_NAN_F;
  }
  return w;
}

// neighbour heights of cell j (0..3) of the owned quad, in reference order k=0..7
__device__ __forceinline__ void gather8(const row6& up, const row6& cu, const row6& dn, int j, float& hc, float nb[8]) {
  const float u[6] = {up.l, up.a, up.b, up.c, up.d, up.r};
  const float c[6] = {cu.l, cu.a, cu.b, cu.c, cu.d, cu.r};
  const float d[6] = {dn.l, dn.a, dn.b, dn.c, dn.d, dn.r};
  hc = c[j + 1];
  nb[0] = u[j + 1];  // (-1, 0)
  nb[1] = c[j];      // ( 0,-1)
  nb[2] = c[j + 2];  // ( 0, 1)
  nb[3] = d[j + 1];  // ( 1, 0)
  nb[4] = u[j];      // (-1,-1)
  nb[5] = u[j + 2];  // (-1, 1)
  nb[6] = d[j];      // ( 1,-1)
  nb[7] = d[j + 2];  // ( 1, 1)
}

__device__ __forceinline__ unsigned valid_mask(int x, int y, int nx, int ny) {
  const bool xm = x > 0, xp = x + 1 < nx, ym = y > 0, yp = y + 1 < ny;
  return (xm ? 1u : 0u) | (ym ? 2u : 0u) | (yp ? 4u : 0u) | (xp ? 8u : 0u) | ((xm && ym) ? 16u : 0u) |
         ((xm && yp) ? 32u : 0u) | ((xp && ym) ? 64u : 0u) | ((xp && yp) ? 128u : 0u);
}

__device__ __forceinline__ int flat_of(int k, int x, int y, int ny) {
  // graph.hpp:35-44 as packed two-bit fields (an indexed constant-bank load serialises over the distinct k of a warp)
  const int dx = (int)((0xA094u >> (2 * k)) & 3u) - 1, dy = (int)((0x8861u >> (2 * k)) & 3u) - 1;
  return (x + dx) * ny + (y + dy);
}

// ---- per-cell functors -----------------------------------------------------

// x / sqrtf(2.0f), correctly rounded, without the generic division sequence.  With c = 0x3FB504F3 and
// r = RN(1/c): q0 = RN(x*r), e = RN(x - q0*c) (exact, one FMA), q = RN(q0 + e*r).  For a fixed divisor this
// Markstein-style correction returns RN(x/c) whenever no intermediate under/overflows; tests/test_gpu_flow.py
// checks it against the IEEE division for ALL 2^32 bit patterns of x.  Tiny and non-finite x take the
// generic division (never on real DEMs).
__device__ __forceinline__ float div_sqrt2(float x) {
  const float c = kSqrt2;
  const float r = 0.707106769f;  // RN(1 / 1.41421354f) = 0x3F3504F3
  const float ax = fabsf(x);
  if (!(ax >= 1.0e-30f && ax <= 1.0e+30f)) return x / c;  // zero, subnormal range, inf, NaN: exact path
  const float q0 = x * r;
  const float e = __fmaf_rn(-q0, c, x);
  return __fmaf_rn(e, r, q0);
}

// the three-instruction quotient alone; valid when x is zero or 1e-30 <= |x| <= 1e30 (see div_sqrt2)
__device__ __forceinline__ float div_sqrt2_fast(float x) {
  const float q0 = x * 0.707106769f;
  const float e = __fmaf_rn(-q0, kSqrt2, x);
  return __fmaf_rn(e, 0.707106769f, q0);
}
// graph.cu:42-60 / :216-233: strict '>' from smax = 0, first maximum wins.  Neighbours outside the map
// arrive as NaN from load_row, and a NaN slope never compares greater - the same cells the reference skips
// with its oob test (a NaN height inside the map behaves identically in the reference: scur is NaN).
template <int K, bool CODE>
struct pick_steepest {
  template <bool EXACT>
  __device__ __forceinline__ int pick(int x, int y, int ny, float hc, const float nb[8], bool& ok) const {
    float smax = 0.0f;
    int next = -1;
    float dh[8];
#pragma unroll
    for (int k = 0; k < K; ++k) dh[k] = hc - nb[k];
    ok = true;
    if (K == 8 && !EXACT) {
      // one range test for the four diagonal differences (NaN and inf fail the upper bound, tiny non-zero
      // values the lower one); the caller redoes the rare failing cell with IEEE divisions
      const unsigned lo = 0x0DA24260u - 1u, hi = 0x7149F2CAu;  // bits of 1e-30f (minus the wrap) and of 1e30f
      const unsigned a4 = __float_as_uint(dh[4]) & 0x7fffffffu, a5 = __float_as_uint(dh[5]) & 0x7fffffffu;
      const unsigned a6 = __float_as_uint(dh[6]) & 0x7fffffffu, a7 = __float_as_uint(dh[7]) & 0x7fffffffu;
      const unsigned top = max(max(a4, a5), max(a6, a7));
      const unsigned bot = min(min(a4 - 1u, a5 - 1u), min(a6 - 1u, a7 - 1u));
      ok = (top <= hi) && (bot >= lo);
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float scur = (k < 4) ? dh[k] : (EXACT ? dh[k] / kSqrt2 : div_sqrt2_fast(dh[k]));
      if (scur > smax) { smax = scur; next = k; }
    }
    if (CODE || next < 0) return next;
    // flat index of the receiver: the shift table (graph.hpp:35-44) packed two bits per entry, biased by one
    const int dx = (int)((0xA094u >> (2 * next)) & 3u) - 1;
    const int dy = (int)((0x8861u >> (2 * next)) & 3u) - 1;
    return (x + dx) * ny + (y + dy);
  }
  __device__ __forceinline__ int operator()(int x, int y, int ny, float hc, const float nb[8], unsigned) const {
    bool ok;
    int r = pick<false>(x, y, ny, hc, nb, ok);
    if (!ok) r = pick<true>(x, y, ny, hc, nb, ok);  // rare: IEEE divisions, everything still in registers
    return r;
  }
};

// curand(): one XORWOW step (curand_kernel.h: curand(curandStateXORWOW_t*))
struct xw_state {
  unsigned v0, v1, v2, v3, v4, d;
};
__device__ __forceinline__ unsigned xw_next(xw_state& s) {
  const unsigned t = s.v0 ^ (s.v0 >> 2);
  s.v0 = s.v1; s.v1 = s.v2; s.v2 = s.v3; s.v3 = s.v4;
  s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
  s.d += 362437u;
  return s.v4 + s.d;
}
// _curand_uniform, curand_uniform.h:69-72
__device__ __forceinline__ float xw_uniform(unsigned x) { return x * 2.3283064e-10f + (2.3283064e-10f / 2.0f); }

// graph.cu:120-171.  Two things are done lazily, with the reference's arithmetic wherever a value is formed:
//  - exp(dE / T) (an IEEE division + ex2.approx) only for downhill neighbours - the others contribute exactly 0;
//  - the test u < CDF_k / Z first as u * Z against CDF_k with a margin of 2^-22 (both roundings are within 2^-24, so
//    outside the margin the comparison is decided); only inside it - or when a NaN / inf is involved, which fails both
//    guards - is the quotient formed.  First k that passes wins, as in the reference's scan.
template <int K>
__device__ __forceinline__ int gibbs_pick(int x, int y, int ny, float hc, const float nb[8], unsigned vm, float T,
                                          float uniform) {
  float CDF[K];
  float Z = 0.0f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    CDF[k] = 0.0f;
    if (!(vm & (1u << k))) continue;
    const float dh = hc - nb[k];
    const float dE = (k < 4) ? dh : div_sqrt2(dh);
    float P = 0.0f;
    if (!(dE <= 0.0f)) P = __expf(dE / T);  // NaN takes the exp path like the reference's (dE <= 0) ? 0 : exp
    CDF[k] = Z + P;
    Z += P;
  }
  const float t = uniform * Z;
  int next = -1;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (next >= 0 || !(vm & (1u << k))) continue;
    const float c = CDF[k];
    bool take;
    if (t < c * (1.0f - 2.3841858e-7f)) take = true;        // 2^-22
    else if (t > c * (1.0f + 2.3841858e-7f)) take = false;
    else take = uniform < (c / Z);
    if (take) next = k;
  }
  return next < 0 ? -1 : flat_of(next, x, y, ny);
}

// ---- random_weighted workspace ---------------------------------------------
// 64-byte header, then six uint32 planes (v0..v4, d) of `pitch` cells each.
struct rw_header {
  unsigned long long magic, seed, next_offset;
  int nx, ny;
  unsigned long long pad[4];
};
static_assert(sizeof(rw_header) == 64, "rw_header");
constexpr unsigned long long kRwMagic = 0x534f494c52573031ull;  // "SOILRW01"

__host__ __device__ inline size_t rw_pitch(long long n) { return (size_t)((n + 63) / 64 * 64); }

struct rw_args {
  const rw_header* hdr;  // null -> stateless (seed every cell like the reference)
  unsigned* planes;
  size_t pitch;
  unsigned long long seed, offset;
  unsigned long long base;  // subsequence of cell 0: the global index of the block's first cell (row strips), else 0
  float T;
};

template <int K>
struct pick_gibbs {
  static constexpr bool kNeedsMask = true;
  rw_args a;
  bool resume;  // header matched: advance the stored streams by one draw
  __device__ __forceinline__ int operator()(int x, int y, int ny, float hc, const float nb[8], unsigned vm) const {
    const size_t n = (size_t)x * ny + y;
    xw_state s;
    if (resume) {
      s.v0 = a.planes[n]; s.v1 = a.planes[a.pitch + n]; s.v2 = a.planes[2 * a.pitch + n];
      s.v3 = a.planes[3 * a.pitch + n]; s.v4 = a.planes[4 * a.pitch + n]; s.d = a.planes[5 * a.pitch + n];
    } else {
      curandStateXORWOW_t st;
      curand_init(a.seed, a.base + (unsigned long long)n, a.offset, &st);  // graph.cu:100
      s.v0 = st.v[0]; s.v1 = st.v[1]; s.v2 = st.v[2]; s.v3 = st.v[3]; s.v4 = st.v[4]; s.d = st.d;
    }
    const float u = xw_uniform(xw_next(s));
    if (a.planes) {
      a.planes[n] = s.v0; a.planes[a.pitch + n] = s.v1; a.planes[2 * a.pitch + n] = s.v2;
      a.planes[3 * a.pitch + n] = s.v3; a.planes[4 * a.pitch + n] = s.v4; a.planes[5 * a.pitch + n] = s.d;
    }
    return gibbs_pick<K>(x, y, ny, hc, nb, vm, a.T, u);
  }
};

// ---- the sweep ---------------------------------------------------------------

template <class F>
__device__ __forceinline__ void sweep_band(const float* __restrict__ h, int32_t* __restrict__ out, int nx, int ny,
                                           int band_rows, const F& f) {
  const int lane = threadIdx.x & 31;
  const int warp_in_block = threadIdx.x >> 5;
  const int strips = (ny + 127) / 128;
  const long long wid = (long long)blockIdx.x * kWarpsPerBlock + warp_in_block;
  const int strip = (int)(wid % strips);
  const int band = (int)(wid / strips);
  const int x0 = band * band_rows;
  if (x0 >= nx) return;
  const int y0 = strip * 128 + lane * 4;
  const bool active = y0 < ny;
  const int x1 = min(x0 + band_rows, nx);

  row6 up = load_row(h, x0 - 1, y0, nx, ny, lane, active);
  row6 cu = load_row(h, x0, y0, nx, ny, lane, active);
  for (int x = x0; x < x1; ++x) {
    const row6 dn = load_row(h, x + 1, y0, nx, ny, lane, active);
    if (active) {
      int4 res;
      int* r = reinterpret_cast<int*>(&res);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float hc, nb[8];
        gather8(up, cu, dn, j, hc, nb);
        r[j] = f(x, y0 + j, ny, hc, nb, F::kNeedsMask ? valid_mask(x, y0 + j, nx, ny) : 0u);
      }
      *reinterpret_cast<int4*>(out + (size_t)x * ny + y0) = res;
    }
    up = cu;
    cu = dn;
  }
}

template <int K>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_sweep_gibbs(const float* __restrict__ h,
                                                                     int32_t* __restrict__ out, int nx, int ny,
                                                                     int band_rows, rw_args a) {
  pick_gibbs<K> f;
  f.a = a;
  f.resume = false;
  if (a.hdr) {
    const rw_header hd = *a.hdr;
    f.resume = hd.magic == kRwMagic && hd.seed == a.seed && hd.next_offset == a.offset && hd.nx == nx && hd.ny == ny &&
               hd.pad[0] == a.base;
  }
  sweep_band(h, out, nx, ny, band_rows, f);
}

// one thread per cell: any row length / alignment
template <class F>
__device__ __forceinline__ void scalar_cell(const float* __restrict__ h, int32_t* __restrict__ out, int nx, int ny,
                                            const F& f) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (long long)nx * ny) return;
  const int x = (int)(n / ny), y = (int)(n % ny);
  const unsigned vm = valid_mask(x, y, nx, ny);
  const float hc = __ldg(h + n);
  float nb[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) nb[k] = (vm & (1u << k)) ? __ldg(h + (size_t)(x + c_shift_x[k]) * ny + (y + c_shift_y[k])) : CUDART_NAN_F;
  out[n] = f(x, y, ny, hc, nb, vm);
}

template <class F>
__global__ void __launch_bounds__(256) k_scalar(const float* __restrict__ h, int32_t* __restrict__ out, int nx, int ny,
                                                F f) {
  scalar_cell(h, out, nx, ny, f);
}

template <int K>
__global__ void __launch_bounds__(256) k_scalar_gibbs(const float* __restrict__ h, int32_t* __restrict__ out, int nx,
                                                      int ny, rw_args a) {
  pick_gibbs<K> f;
  f.a = a;
  f.resume = false;
  if (a.hdr) {
    const rw_header hd = *a.hdr;
    f.resume = hd.magic == kRwMagic && hd.seed == a.seed && hd.next_offset == a.offset && hd.nx == nx && hd.ny == ny &&
               hd.pad[0] == a.base;
  }
  scalar_cell(h, out, nx, ny, f);
}

// (Re)seeds the per-cell streams in the workspace unless its header says they already stand at (seed, offset).
// curand_init(seed, n, offset) costs ~20 matrix-vector products of the 160-bit state per cell (skip-ahead by
// n * 2^67); the state of cell n + 16 is the state of cell n skipped ahead by 16 subsequences - ONE product
// (16 = 4^2, a single precomputed matrix).  So a thread seeds cell n0 in full and then walks n0 + 16 i; the 16
// threads of a half-warp store 64 contiguous bytes per plane.  Same states, bit for bit, ~10x cheaper.
constexpr int kSeedChain = 64;
__global__ void __launch_bounds__(128) k_rw_seed(const rw_header* hdr, unsigned* __restrict__ planes, size_t pitch,
                                                 long long N, unsigned long long seed, unsigned long long offset,
                                                 unsigned long long base, int nx, int ny) {
  const rw_header hd = *hdr;
  if (hd.magic == kRwMagic && hd.seed == seed && hd.next_offset == offset && hd.nx == nx && hd.ny == ny &&
      hd.pad[0] == base)
    return;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n0 = (t >> 4) * (16LL * kSeedChain) + (t & 15);
  if (n0 >= N) return;
  curandStateXORWOW_t st;
  curand_init(seed, base + (unsigned long long)n0, offset, &st);  // graph.cu:100
  for (int i = 0; i < kSeedChain; ++i) {
    const long long n = n0 + 16LL * i;
    if (n >= N) break;
    planes[n] = st.v[0]; planes[pitch + n] = st.v[1]; planes[2 * pitch + n] = st.v[2];
    planes[3 * pitch + n] = st.v[3]; planes[4 * pitch + n] = st.v[4]; planes[5 * pitch + n] = st.d;
    skipahead_sequence(16ull, &st);
  }
}

__global__ void k_rw_header(rw_header* hdr, unsigned long long seed, unsigned long long next_offset,
                            unsigned long long base, int nx, int ny) {
  hdr->pad[0] = base;
  hdr->magic = kRwMagic;
  hdr->seed = seed;
  hdr->next_offset = next_offset;
  hdr->nx = nx;
  hdr->ny = ny;
}

__global__ void __launch_bounds__(256) k_seed(curandStateXORWOW_t* st, long long n, unsigned long long seed,
                                              unsigned long long first, unsigned long long offset) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  curandStateXORWOW_t s;
  curand_init(seed, first + (unsigned long long)i, offset, &s);
  st[i] = s;
}

// ---- steepest / direction: edge-shared sweep ---------------------------------
// Every D8 edge joins two cells, and the two cells see slopes of opposite sign along it:
// hA - hB == -(hB - hA) exactly, and x / sqrt2 is odd in x under round-to-nearest.  So a lane computes each
// difference and each diagonal quotient ONCE (the edges from the current row down to the next one and along
// the current row) and carries the downward edges to the next iteration, where they are the upward edges
// with the sign flipped: 19 subtractions and 10 quotients per 4 cells instead of 32 and 16.
//
// The quotients use the three-instruction div_sqrt2_fast.  It differs from the IEEE division only for
// 0 < |x| <= 2.2e-32, where both results stay below 1.6e-32 and keep the sign of x, and for x = +-inf
// (NaN instead of inf) - soilb200_selftest_div_sqrt2 re-proves this over all 2^32 inputs.  Hence the winner
// of the scan is the reference's winner whenever the winning slope is 0 (nothing positive) or >= 1e-30 and no
// height of the 3x3 window exceeds 1e38 in magnitude (no difference overflows); any other quad is redone from
// memory with IEEE divisions (exact_cell).  Cells outside the map are NaN in the window: NaN slopes never
// compare greater, like the reference's oob skip.

// graph.cu:42-60 / :216-233 verbatim arithmetic for one cell, straight from memory
template <int K, bool CODE>
__device__ __noinline__ int exact_cell(const float* __restrict__ h, int x, int y, int nx, int ny) {
  const unsigned vm = valid_mask(x, y, nx, ny);
  const float hc = __ldg(h + (size_t)x * ny + y);
  float smax = 0.0f;
  int next = -1;
  for (int k = 0; k < K; ++k) {
    if (!(vm & (1u << k))) continue;
    const float dh = hc - __ldg(h + (size_t)(x + c_shift_x[k]) * ny + (y + c_shift_y[k]));
    const float scur = (k < 4) ? dh : dh / kSqrt2;
    if (scur > smax) { smax = scur; next = k; }
  }
  if (CODE || next < 0) return next;
  return flat_of(next, x, y, ny);
}

struct win6 {
  float w[6];  // columns y0-1 .. y0+4 of one row
  bool big;    // some |height| > 1e38
};

__device__ __forceinline__ win6 load_win(const float* __restrict__ h, int x, int y0, int nx, int ny, bool okL,
                                         bool okR) {
  win6 r;
#pragma unroll
  for (int i = 0; i < 6; ++i) r.w[i] = CUDART_NAN_F;
  r.big = false;
  if (x >= 0 && x < nx) {
    const float* p = h + (size_t)x * ny + y0;
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    r.w[1] = v.x; r.w[2] = v.y; r.w[3] = v.z; r.w[4] = v.w;
    if (okL) r.w[0] = __ldg(p - 1);  // the neighbouring lane's line: an L1 hit
    if (okR) r.w[5] = __ldg(p + 4);
    // fmaxf drops NaN operands, so the out-of-map markers do not trip the flag
    const float m = fmaxf(fmaxf(fmaxf(fabsf(r.w[0]), fabsf(r.w[1])), fmaxf(fabsf(r.w[2]), fabsf(r.w[3]))),
                          fmaxf(fabsf(r.w[4]), fabsf(r.w[5])));
    r.big = m > 1.0e38f;
  }
  return r;
}

// edges from row `a` down to row `b`: vertical d0[j] = a[j]-b[j]; dr[i] = (a[i-1] - b[i]) / sqrt2 for window
// columns i-1 = -1..3 (towards +y); dl[i] = (a[i] - b[i-1]) / sqrt2 for columns i = 0..4 (towards -y)
template <int K>
struct down_edges {
  float d0[4], dr[5], dl[5];
  __device__ __forceinline__ void compute(const win6& a, const win6& b) {
#pragma unroll
    for (int j = 0; j < 4; ++j) d0[j] = a.w[j + 1] - b.w[j + 1];
    if (K == 8) {
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        dr[i] = div_sqrt2_fast(a.w[i] - b.w[i + 1]);
        dl[i] = div_sqrt2_fast(a.w[i + 1] - b.w[i]);
      }
    }
  }
};

// one row of 4 cells: edges down to `dn`, the scan in reference order, the exact redo, the 16-byte store;
// then the window slides (up <- edges just computed, cu <- dn)
template <int K, bool CODE>
__device__ __forceinline__ void steepest_row(const float* __restrict__ h, int32_t* __restrict__ out, int nx, int ny,
                                             int x, int y0, const int (&off)[8], win6& cu, down_edges<K>& up,
                                             bool& bigU, const win6& dn, bool redo = false) {
  down_edges<K> dw;
  dw.compute(cu, dn);
  float hz[5];  // along the row: hz[i] = w[i] - w[i+1]
#pragma unroll
  for (int i = 0; i < 5; ++i) hz[i] = cu.w[i] - cu.w[i + 1];

  int4 res;
  int* r = reinterpret_cast<int*>(&res);
  bool ok = !(bigU || cu.big || dn.big || redo);
  const int rowbase = x * ny + y0;
  // The scan "strict > from 0, first maximum wins" as: the maximum m of {0, slopes} (three-input max, NaN dropped),
  // then the lowest k whose slope equals m, provided m > 0.  Same winner as the running compare - a slope that
  // never beat the running maximum cannot equal the final one unless an earlier k holds the same value - in 21
  // instead of 24 ALU-pipe instructions per cell; this kernel is bound by that pipe.
  unsigned lowest = 0xffffffffu;  // min over the quad of bits(m) - 1, with m == 0 wrapped to the top
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float sl[8] = {-up.d0[j], -hz[j], hz[j + 1], dw.d0[j], -up.dr[j], -up.dl[j + 1], dw.dl[j], dw.dr[j + 1]};
    float m = fmaxf(fmaxf(fmaxf(fmaxf(0.0f, sl[0]), sl[1]), sl[2]), sl[3]);
    if (K == 8) m = fmaxf(fmaxf(fmaxf(fmaxf(m, sl[4]), sl[5]), sl[6]), sl[7]);
    const bool pos = m > 0.0f;
    const int n = rowbase + j;
    int next = CODE ? -1 : ~n;  // steepest: tracked as the offset shift_k, "none" as ~n so that n + offset == -1
    // two levels: the first pair {2g, 2g+1} whose running maximum already equals m, then the member of that pair
    const float g0 = fmaxf(fmaxf(0.0f, sl[0]), sl[1]), g1 = fmaxf(fmaxf(g0, sl[2]), sl[3]);
    const float g2 = (K == 8) ? fmaxf(fmaxf(g1, sl[4]), sl[5]) : m;
    auto pick2 = [&](int k) { return sl[k] == m ? (CODE ? k : off[k]) : (CODE ? k + 1 : off[k + 1]); };
    int win = (K == 8) ? pick2(6) : pick2(2);
    if (K == 8) win = g2 == m ? pick2(4) : win;
    if (K == 8) win = g1 == m ? pick2(2) : win;
    win = g0 == m ? pick2(0) : win;
    next = pos ? win : next;
    lowest = min(lowest, __float_as_uint(m) - 1u);
    r[j] = CODE ? next : n + next;
  }
  ok = ok && lowest >= 0x0DA24260u - 1u;  // every m is 0 or >= 1e-30f
  if (K == 8 && !ok) {
    for (int j = 0; j < 4; ++j) r[j] = exact_cell<K, CODE>(h, x, y0 + j, nx, ny);
  }
  *reinterpret_cast<int4*>(out + rowbase) = res;  // nx * ny < 2^31 (launch_steepest), so the flat index fits an int
  up = dw;
  bigU = cu.big;
  cu = dn;
}

// what load_win derives from a row, for a row that arrived through the shared-memory ring: a row outside the map
// becomes NaN markers, `big` flags a height above 1e38 in magnitude (fmaxf drops NaN operands, so the markers do
// not trip it)
__device__ __forceinline__ void settle_win(win6& r, bool in_map) {
#pragma unroll
  for (int i = 0; i < 6; ++i) r.w[i] = in_map ? r.w[i] : CUDART_NAN_F;
  const float m = fmaxf(fmaxf(fmaxf(fabsf(r.w[0]), fabsf(r.w[1])), fmaxf(fabsf(r.w[2]), fabsf(r.w[3]))),
                        fmaxf(fabsf(r.w[4]), fabsf(r.w[5])));
  r.big = m > 1.0e38f;
}

constexpr int kRing = 6;               // ring slots per warp: 5 rows (2.5 KB) in flight per warp
constexpr int kSteepestBlocksPerSM = 6;  // caps the kernel at 80 registers: 24 warps per SM
constexpr int kRingRow = 4 + 128 + 4;  // [.. left halo | 128 columns | right halo ..], 16-byte granules
template <int K, bool CODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, kSteepestBlocksPerSM) k_steepest(const float* __restrict__ h,
                                                                                        int32_t* __restrict__ out, int nx,
                                                                                        int ny, int band_rows) {
  const int lane = threadIdx.x & 31;
  const int strips = (ny + 127) / 128;
  const long long wid = (long long)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  const int strip = (int)(wid % strips);
  const int x0 = (int)(wid / strips) * band_rows;
  const int y0 = strip * 128 + lane * 4;
  const unsigned lanes = __ballot_sync(0xffffffffu, x0 < nx && y0 < ny);  // the lanes that sweep; the others leave
  if (x0 >= nx || y0 >= ny) return;
  const int x1 = min(x0 + band_rows, nx);
  const bool okL = y0 > 0, okR = y0 + 4 < ny;

  const int off[8] = {-ny, -1, 1, ny, -ny - 1, -ny + 1, ny - 1, ny + 1};  // flat offsets of graph.hpp:35-44

  win6 cu = load_win(h, x0, y0, nx, ny, okL, okR);
  down_edges<K> up;  // edges from row x-1 down to row x
  bool bigU;
  {
    const win6 above = load_win(h, x0 - 1, y0, nx, ny, okL, okR);
    up.compute(above, cu);
    bigU = above.big;
  }
  // Rows travel through a private shared-memory ring per warp: kRing slots of one 128-column row (+ one halo cell
  // each side), filled by 16-byte cp.async copies (each lane its own 4 cells; lanes 0 / 31 also the halo cell) that
  // are issued kRing - 1 rows ahead of their use.  The copies keep kRing - 1 rows per warp in flight without holding
  // registers, which is what the bandwidth needs (bytes in flight = bandwidth x latency ~ 22 KB of loads per SM;
  // with loads into registers one or two rows ahead the warp still waited on its own row, long_scoreboard 3.3
  // cycles per issue).  No block-level barrier: a warp waits for its own copy group and __syncwarp()s before
  // reading its neighbours' cells; a slot is refilled only after the __syncwarp() of the following iteration.
  // Below the last row of the map (D8) `dn` holds a copy of that row instead of the NaN markers: that row is
  // redone by exact_cell, which tests the map bounds itself, and nothing follows the last row.
  __shared__ __align__(16) float ring[kWarpsPerBlock][kRing][kRingRow];
  float (*my)[kRingRow] = ring[threadIdx.x >> 5];
  const int xlast = min(x1, nx - 1);
  const unsigned ring0 = (unsigned)__cvta_generic_to_shared(&my[0][4 + 4 * lane]);  // this lane's cells in slot 0
  const bool haloL = lane == 0 && okL, haloR = lane == 31 && okR;
  const float* src0 = h + y0;
  auto issue = [&](int x, int slot) {
    // rows past the band's lower halo or past the map: the last row the band reads anyway (an L2 hit whose values
    // are never used) - measured 4 % faster than skipping the copies under a branch
    const float* src = src0 + min(x, xlast) * ny;  // nx * ny < 2^31 (launch_steepest): 32-bit index arithmetic
    const unsigned dst = ring0 + (unsigned)(slot * kRingRow * sizeof(float));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
    if (haloL) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst - 4u), "l"(src - 1) : "memory");
    if (haloR) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + 16u), "l"(src + 4) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int d = 0; d < kRing - 1; ++d) issue(x0 + 1 + d, d);
  for (int xb = x0; xb < x1; xb += kRing) {
#pragma unroll
    for (int slot = 0; slot < kRing; ++slot) {  // the slot of row xb + slot + 1 is a compile-time constant
      const int x = xb + slot;
      if (x >= x1) break;
      asm volatile("cp.async.wait_group %0;" ::"n"(kRing - 2) : "memory");
      __syncwarp(lanes);
      win6 dn;
      {
        const float* row = &my[slot][4 + 4 * lane];
        const float4 v = *reinterpret_cast<const float4*>(row);
        dn.w[0] = okL ? row[-1] : CUDART_NAN_F;
        dn.w[1] = v.x; dn.w[2] = v.y; dn.w[3] = v.z; dn.w[4] = v.w;
        dn.w[5] = okR ? row[4] : CUDART_NAN_F;
      }
      issue(x + kRing, (slot + kRing - 1) % kRing);  // the slot read one iteration ago: every lane is past that read
      settle_win(dn, K == 8 || x + 1 < nx);
      steepest_row<K, CODE>(h, out, nx, ny, x, y0, off, cu, up, bigU, dn, K == 8 && x + 1 >= nx);
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");  // nothing may land in the ring after the block is gone
}

// ---- the same sweep with the rows staged through shared memory by the TMA engine -----------------------------
// k_steepest leaves a third of the issue slots idle waiting for its own row (long_scoreboard).  Here a producer
// warp streams the rows of the block's 512-column strip into a ring of kStages shared-memory rows with
// cp.async.bulk (one elected lane; completion counted in bytes on an mbarrier per stage), kStages rows ahead of
// the four consumer warps, which read their 6 heights per row at shared-memory latency and hand the stage back
// through a second mbarrier.  Every row is read from shared memory once (the window above lives in registers).
constexpr int kStagedWarps = 4;                 // consumer warps; + 1 producer warp
constexpr int kStagedCols = kStagedWarps * 128; // columns per block
constexpr int kStages = 8;
constexpr int kStagedBand = 64;                 // rows per block
constexpr int kStageFloats = 4 + kStagedCols + 4;  // [4 left halo | 512 | 4 right halo], 16-byte granules

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned done = 0, spins = 0;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (!done && ++spins > (1u << 22)) __trap();  // a lost arrival must fail the launch, never hang the GPU
  } while (!done);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

template <int K, bool CODE>
__global__ void __launch_bounds__((kStagedWarps + 1) * 32) k_steepest_staged(const float* __restrict__ h,
                                                                             int32_t* __restrict__ out, int nx,
                                                                             int ny) {
  __shared__ __align__(128) float ring[kStages][kStageFloats];
  __shared__ __align__(8) unsigned long long full[kStages], empty[kStages];
  const int strips = (ny + kStagedCols - 1) / kStagedCols;
  const int yb = (int)(blockIdx.x % strips) * kStagedCols;  // first column of the block
  const int x0 = (int)(blockIdx.x / strips) * kStagedBand;
  const int x1 = min(x0 + kStagedBand, nx);
  const int cols = min(kStagedCols, ny - yb);  // multiple of 4
  const int rows = (x1 - x0) + 2;              // map rows x0-1 .. x1 pass through the ring
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);              // the producer's arrive (+ the bytes of the row)
      mbar_init(&empty[s], kStagedWarps);  // one arrive per consumer warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kStagedWarps) {  // ---- producer: one lane feeds the ring
    if (lane != 0) return;
    const bool haveL = yb > 0, haveR = yb + cols < ny;
    const unsigned bytes = (unsigned)cols * 4u + (haveL ? 16u : 0u) + (haveR ? 16u : 0u);
    for (int r = 0; r < rows; ++r) {
      const int s = r % kStages;
      if (r >= kStages) mbar_wait(&empty[s], (unsigned)((r / kStages - 1) & 1));
      const int x = x0 - 1 + r;
      if (x < 0 || x >= nx) {  // outside the map: nothing to copy, the consumers fill NaN themselves
        mbar_arrive(&full[s]);
        continue;
      }
      const float* src = h + (size_t)x * ny + yb;
      mbar_arrive_expect_tx(&full[s], bytes);
      bulk_g2s(&ring[s][4], src, (unsigned)cols * 4u, &full[s]);
      if (haveL) bulk_g2s(&ring[s][0], src - 4, 16u, &full[s]);
      if (haveR) bulk_g2s(&ring[s][4 + cols], src + cols, 16u, &full[s]);
    }
    return;
  }

  // ---- consumers: warp w owns columns yb + 128 w .. +127, lane 4 of them
  const int c = warp * 128 + lane * 4;  // column within the block
  const int y0 = yb + c;
  const bool active = c < cols;
  const bool okL = y0 > 0, okR = y0 + 4 < ny;
  const int off[8] = {-ny, -1, 1, ny, -ny - 1, -ny + 1, ny - 1, ny + 1};  // flat offsets of graph.hpp:35-44

  // row r of the stream out of stage s (compile-time after unrolling), then the stage goes back to the producer
  auto take = [&](int r, int s, unsigned parity) -> win6 {
    mbar_wait(&full[s], parity);
    win6 w;
#pragma unroll
    for (int i = 0; i < 6; ++i) w.w[i] = CUDART_NAN_F;
    w.big = false;
    const int x = x0 - 1 + r;
    if (active && x >= 0 && x < nx) {
      const float4 v = *reinterpret_cast<const float4*>(&ring[s][4 + c]);
      w.w[1] = v.x; w.w[2] = v.y; w.w[3] = v.z; w.w[4] = v.w;
      if (okL) w.w[0] = ring[s][4 + c - 1];
      if (okR) w.w[5] = ring[s][4 + c + 4];
      const float m = fmaxf(fmaxf(fmaxf(fabsf(w.w[0]), fabsf(w.w[1])), fmaxf(fabsf(w.w[2]), fabsf(w.w[3]))),
                            fmaxf(fabsf(w.w[4]), fabsf(w.w[5])));
      w.big = m > 1.0e38f;
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
    return w;
  };

  win6 cu;
  down_edges<K> up;
  bool bigU = false;
  unsigned parity = 0;
  for (int base = 0; base < rows; base += kStages, parity ^= 1u) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      const int r = base + s;
      if (r >= rows) break;
      const win6 dn = take(r, s, parity);
      if (r == 0) {
        cu = dn;  // the row above the band
      } else if (r == 1) {
        up.compute(cu, dn);
        bigU = cu.big;
        cu = dn;
      } else if (active) {
        steepest_row<K, CODE>(h, out, nx, ny, x0 + r - 2, y0, off, cu, up, bigU, dn);
      }
    }
  }
}

// ---- the same sweep fed by 2-D tensor-map loads (TMA tile mode) ------------------------------------------------
// A producer lane issues cp.async.bulk.tensor.2d loads of kTmaRows x 256-column boxes of the height map (two per
// stage for the block's 512 columns, plus a 4-column box on either side for the halo cells) through a tensor map
// over THIS GPU's height array; completion is counted in bytes on an mbarrier per stage.  Two differences from the
// 1-D staged kernel above: a stage holds several rows, so the producer / consumer hand-shake is paid once per
// kTmaRows rows instead of once per row, and the tensor map fills what lies outside the map with NaN (out-of-bound
// fill), which is exactly the marker the sweep uses for "no neighbour" (NaN never wins the strict > scan,
// graph.cu:42-60) - the rows above / below the map and the columns beside it need no special case at all.
constexpr int kTmaRows = 4;        // rows per stage
constexpr int kTmaStages = 4;
constexpr int kTmaBox = 256;       // columns per box: the largest a tensor map's box may be
constexpr int kTmaBand = 64;       // rows per block

struct __align__(128) tma_stage {     // every box lands on a 128-byte boundary
  float main[2][kTmaRows][kTmaBox];  // columns yb .. yb+255 and yb+256 .. yb+511
  float left[kTmaRows][4];           // columns yb-4 .. yb-1
  float pad_l[32 - kTmaRows * 4];
  float right[kTmaRows][4];          // columns yb+512 .. yb+515
  float pad_r[32 - kTmaRows * 4];
};
static_assert(sizeof(tma_stage) % 128 == 0 && kTmaRows * 4 < 32, "tma_stage");
constexpr unsigned kTmaStageBytes = 2u * kTmaRows * kTmaBox * 4u + 2u * kTmaRows * 16u;

__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

template <int K, bool CODE>
__global__ void __launch_bounds__((kStagedWarps + 1) * 32) k_steepest_tma(const __grid_constant__ CUtensorMap map_main,
                                                                          const __grid_constant__ CUtensorMap map_halo,
                                                                          const float* __restrict__ h,
                                                                          int32_t* __restrict__ out, int nx, int ny) {
  extern __shared__ __align__(128) unsigned char tma_smem[];
  tma_stage* ring = reinterpret_cast<tma_stage*>(tma_smem);
  __shared__ __align__(8) unsigned long long full[kTmaStages], empty[kTmaStages];
  const int strips = (ny + kStagedCols - 1) / kStagedCols;
  const int yb = (int)(blockIdx.x % strips) * kStagedCols;  // first column of the block
  const int x0 = (int)(blockIdx.x / strips) * kTmaBand;
  const int x1 = min(x0 + kTmaBand, nx);
  const int rows = (x1 - x0) + 2;                             // map rows x0-1 .. x1 pass through the ring
  const int n_stages = (rows + kTmaRows - 1) / kTmaRows;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTmaStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kStagedWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kStagedWarps) {  // ---- producer: one lane, four box loads per stage
    if (lane != 0) return;
    for (int g = 0; g < n_stages; ++g) {
      const int s = g % kTmaStages;
      if (g >= kTmaStages) mbar_wait(&empty[s], (unsigned)((g / kTmaStages - 1) & 1));
      const int xr = x0 - 1 + g * kTmaRows;                  // may be -1, may run past the map: filled with NaN
      mbar_arrive_expect_tx(&full[s], kTmaStageBytes);
      tma_load_2d(&ring[s].main[0][0][0], &map_main, yb, xr, &full[s]);
      tma_load_2d(&ring[s].main[1][0][0], &map_main, yb + kTmaBox, xr, &full[s]);
      tma_load_2d(&ring[s].left[0][0], &map_halo, yb - 4, xr, &full[s]);
      tma_load_2d(&ring[s].right[0][0], &map_halo, yb + 2 * kTmaBox, xr, &full[s]);
    }
    return;
  }

  // ---- consumers: warp w owns columns yb + 128 w .. +127, lane 4 of them
  const int c = warp * 128 + lane * 4;  // column within the block
  const int y0 = yb + c;
  const bool active = y0 < ny;
  const int half = c >> 8, cc = c & (kTmaBox - 1);
  const int off[8] = {-ny, -1, 1, ny, -ny - 1, -ny + 1, ny - 1, ny + 1};  // flat offsets of graph.hpp:35-44

  win6 cu;
  down_edges<K> up;
  bool bigU = false;
  for (int g = 0; g < n_stages; ++g) {
    const int s = g % kTmaStages;
    mbar_wait(&full[s], (unsigned)((g / kTmaStages) & 1));
    const tma_stage& st = ring[s];
#pragma unroll
    for (int i = 0; i < kTmaRows; ++i) {
      const int r = g * kTmaRows + i;
      if (r >= rows) break;
      win6 dn;
      const float4 v = *reinterpret_cast<const float4*>(&st.main[half][i][cc]);
      dn.w[1] = v.x; dn.w[2] = v.y; dn.w[3] = v.z; dn.w[4] = v.w;
      dn.w[0] = c == 0 ? st.left[i][3] : (cc == 0 ? st.main[0][i][kTmaBox - 1] : st.main[half][i][cc - 1]);
      dn.w[5] = c + 4 == 2 * kTmaBox ? st.right[i][0] : (cc + 4 == kTmaBox ? st.main[1][i][0] : st.main[half][i][cc + 4]);
      const float m = fmaxf(fmaxf(fmaxf(fabsf(dn.w[0]), fabsf(dn.w[1])), fmaxf(fabsf(dn.w[2]), fabsf(dn.w[3]))),
                            fmaxf(fabsf(dn.w[4]), fabsf(dn.w[5])));
      dn.big = m > 1.0e38f;
      if (r == 0) {
        cu = dn;  // the row above the band
      } else if (r == 1) {
        up.compute(cu, dn);
        bigU = cu.big;
        cu = dn;
      } else if (active) {
        steepest_row<K, CODE>(h, out, nx, ny, x0 + r - 2, y0, off, cu, up, bigU, dn);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
  }
}

// self-test: every binary32 bit pattern through div_sqrt2 against the IEEE division
__global__ void __launch_bounds__(256) k_selftest_div_sqrt2(unsigned long long* mismatches) {
  unsigned long long bad = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32);
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const float x = __uint_as_float((unsigned)i);
    const float exact = x / kSqrt2;
    const unsigned a = __float_as_uint(div_sqrt2(x)), b = __float_as_uint(exact);
    const bool both_nan = (x != x);
    if (a != b && !both_nan) ++bad;
    // what k_steepest relies on: the unguarded quotient is exact from 1e-30 up to FLT_MAX, and below that
    // both results stay under 1e-30 with the sign of x
    const float fast = div_sqrt2_fast(x);
    const float ax = fabsf(x);
    if (ax >= 1.0e-30f && ax <= 3.402823466e+38f) {
      if (__float_as_uint(fast) != b) ++bad;
    } else if (ax < 1.0e-30f) {
      if (!(fabsf(fast) < 1.0e-30f) || !(fabsf(exact) < 1.0e-30f) || (x > 0.0f && !(fast > 0.0f)) ||
          (x < 0.0f && !(fast < 0.0f)))
        ++bad;
    }
  }
  if (bad) atomicAdd(mismatches, bad);
}

inline bool vector_ok(const void* a, const void* b, int ny) {
  return (ny % 4 == 0) && ((reinterpret_cast<uintptr_t>(a) & 15) == 0) && ((reinterpret_cast<uintptr_t>(b) & 15) == 0);
}

// rows per warp.  Small maps: few rows, so that the grid still fills the GPU (a 1024^2 map in 32-row bands is 256
// warps on 148 SMs).  Large maps: a launch runs in waves of sm_count x resident blocks, every block taking about
// (rows + halo) row-times, so the band height is the one in [kBandRows/2, 3 kBandRows/2] that minimises
// waves x (rows + halo) - at 8192^2 with 8 resident blocks per SM, 37-row bands fill 3 waves exactly where 32-row
// bands leave the fourth wave half empty.
template <class Kernel>
inline int resident_blocks(Kernel k, int threads) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k, threads, 0) != cudaSuccess || n < 1) { cudaGetLastError(); n = 1; }
  return n;
}
inline int sweep_band_rows(int nx, int ny, int blocks_per_sm, int halo) {
  const long long strips = (ny + 127) / 128;
  const long long want = (long long)sm_count() * 48;  // warps
  long long rows = (long long)nx * strips / want;
  if (rows < 4) rows = 4;
  if (rows < kBandRows / 2) return (int)rows;
  const long long capacity = (long long)sm_count() * blocks_per_sm;
  long long best = 0, best_cost = 0;
  for (long long r = kBandRows / 2; r <= kBandRows * 3 / 2; ++r) {
    const long long bands = (nx + r - 1) / r, blocks = (strips * bands + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const long long cost = ((blocks + capacity - 1) / capacity) * (r + halo);
    if (!best || cost < best_cost || (cost == best_cost && r > best)) { best = r; best_cost = cost; }
  }
  return (int)best;
}
inline unsigned sweep_blocks(int nx, int ny, int band_rows) {
  const long long strips = (ny + 127) / 128, bands = (nx + band_rows - 1) / band_rows;
  return (unsigned)((strips * bands + kWarpsPerBlock - 1) / kWarpsPerBlock);
}

// The register-window sweep is the default: measured at 8192^2 the staged kernel hides the row latency
// (long_scoreboard 5.5 -> 0.9 cycles per issue, issue slots 67 -> 76 % busy) but pays 35 % more instructions for the
// ring (mbarrier waits and arrivals, producer warp) and ends 12 % slower - the sweep is issue-bound, not latency-
// bound.  SOILB200_STEEPEST=staged selects it (tests/test_gpu_flow.py runs both).
inline bool use_staged_sweep(int, int) {
  static const bool staged = [] {
    const char* e = std::getenv("SOILB200_STEEPEST");
    return e && e[0] == 's';
  }();
  return staged;
}
inline bool use_tma_sweep() {  // SOILB200_STEEPEST=tma: rows arrive as 2-D tensor-map boxes (k_steepest_tma)
  static const bool tma = [] {
    const char* e = std::getenv("SOILB200_STEEPEST");
    return e && e[0] == 't';
  }();
  return tma;
}

// tensor map over the (nx, ny) float map at `h` (this GPU's memory) with boxes of box_cols x kTmaRows and NaN as the
// out-of-bound fill; false if the driver entry point is missing or refuses the shape
inline bool make_height_map(CUtensorMap* map, const float* h, int nx, int ny, int box_cols) {
  typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static encode_fn encode = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      fn = nullptr;
    return (encode_fn)fn;
  }();
  if (!encode) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)ny, (cuuint64_t)nx};        // innermost first
  const cuuint64_t strides[1] = {(cuuint64_t)ny * sizeof(float)};     // bytes between rows (a multiple of 16: ny % 4 == 0)
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)kTmaRows};
  const cuuint32_t estr[2] = {1, 1};
  return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(h), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA) == CUDA_SUCCESS;
}

template <bool CODE>
int launch_steepest(const float* h, int32_t* out, int nx, int ny, int edge, cudaStream_t s) {
  if (!h || !out || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if (edge != SOILB200_D4 && edge != SOILB200_D8) return SOILB200_E_EDGE;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  const long long N = (long long)nx * ny;
  if (vector_ok(h, out, ny)) {
    static const int resident = resident_blocks(k_steepest<8, CODE>, kWarpsPerBlock * 32);
    const int br = sweep_band_rows(nx, ny, resident, 2);
    CUtensorMap map_main, map_halo;
    if (use_tma_sweep() && make_height_map(&map_main, h, nx, ny, kTmaBox) && make_height_map(&map_halo, h, nx, ny, 4)) {
      const unsigned nb = (unsigned)(((ny + kStagedCols - 1) / kStagedCols) * (long long)((nx + kTmaBand - 1) / kTmaBand));
      const size_t smem = sizeof(tma_stage) * kTmaStages;
      if (edge == SOILB200_D4) k_steepest_tma<4, CODE><<<nb, (kStagedWarps + 1) * 32, smem, s>>>(map_main, map_halo, h, out, nx, ny);
      else k_steepest_tma<8, CODE><<<nb, (kStagedWarps + 1) * 32, smem, s>>>(map_main, map_halo, h, out, nx, ny);
    } else if (use_staged_sweep(nx, ny)) {  // opt-in: rows staged through shared memory by bulk async copies
      const unsigned nb = (unsigned)(((ny + kStagedCols - 1) / kStagedCols) * (long long)((nx + kStagedBand - 1) / kStagedBand));
      if (edge == SOILB200_D4) k_steepest_staged<4, CODE><<<nb, (kStagedWarps + 1) * 32, 0, s>>>(h, out, nx, ny);
      else k_steepest_staged<8, CODE><<<nb, (kStagedWarps + 1) * 32, 0, s>>>(h, out, nx, ny);
    } else if (edge == SOILB200_D4) k_steepest<4, CODE><<<sweep_blocks(nx, ny, br), kWarpsPerBlock * 32, 0, s>>>(h, out, nx, ny, br);
    else k_steepest<8, CODE><<<sweep_blocks(nx, ny, br), kWarpsPerBlock * 32, 0, s>>>(h, out, nx, ny, br);
  } else {
    if (edge == SOILB200_D4) k_scalar<<<blocks_for(N, 256), 256, 0, s>>>(h, out, nx, ny, pick_steepest<4, CODE>());
    else k_scalar<<<blocks_for(N, 256), 256, 0, s>>>(h, out, nx, ny, pick_steepest<8, CODE>());
  }
  count_launch();
  return check_launch();
}

// ---- normal map ---------------------------------------------------------------
// SURVEY.md App. A-8 gradient (the stencil body lives in the absent silt; this is our stated definition): per
// axis the 4th-order central difference ((f-2 - 8 f-1) + (8 f+1 - f+2)) * (1/12) when all four neighbours exist
// (in bounds and not NaN), else the 2nd-order central one, else one-sided, else 0; then scaled by the
// precomputed ratios scale.z/scale.x and scale.z/scale.y.  op/normal.hpp:31-36: normalize(vec3(-g.x, -g.y, 1)).
constexpr float kTwelfth = 1.0f / 12.0f;

__device__ __forceinline__ float axis_grad(float fm2, float fm1, float f0, float fp1, float fp2) {
  const bool m2 = fm2 == fm2, m1 = fm1 == fm1, p1 = fp1 == fp1, p2 = fp2 == fp2;
  if (f0 != f0) return 0.0f;
  if (m2 && m1 && p1 && p2) return ((fm2 - 8.0f * fm1) + (8.0f * fp1 - fp2)) * kTwelfth;
  if (m1 && p1) return (fp1 - fm1) * 0.5f;
  if (p1) return fp1 - f0;
  if (m1) return f0 - fm1;
  return 0.0f;
}

__device__ __forceinline__ void unit_normal(float dx, float dy, float rx, float ry, float& ox, float& oy, float& oz) {
  const float vx = -(dx * rx), vy = -(dy * ry);
  const float inv = 1.0f / sqrtf(vx * vx + vy * vy + 1.0f);  // glm::normalize = v * inversesqrt(dot(v, v))
  ox = vx * inv;
  oy = vy * inv;
  oz = inv;
}

// one thread per cell: any row length / alignment
__global__ void __launch_bounds__(256) k_normal(const float* __restrict__ h, float* __restrict__ out, int nx, int ny,
                                                float rx, float ry) {
  const int y = blockIdx.x * 32 + (threadIdx.x & 31);
  const int x = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= nx || y >= ny) return;
  auto at = [&](int xx, int yy) -> float {
    return (xx < 0 || yy < 0 || xx >= nx || yy >= ny) ? CUDART_NAN_F : __ldg(h + (size_t)xx * ny + yy);
  };
  const float f0 = at(x, y);
  const float dx = axis_grad(at(x - 2, y), at(x - 1, y), f0, at(x + 1, y), at(x + 2, y));
  const float dy = axis_grad(at(x, y - 2), at(x, y - 1), f0, at(x, y + 1), at(x, y + 2));
  float* o = out + 3 * ((size_t)x * ny + y);
  unit_normal(dx, dy, rx, ry, o[0], o[1], o[2]);
}

// the sweep: a lane owns 4 columns and keeps rows x-2..x+2 of them in registers; the +-2 column taps of the
// current row are two 8-byte loads that hit the neighbouring lanes' lines; the 12 output floats of a lane are
// contiguous and leave as three 16-byte stores.  The 4th-order formula propagates NaN, so a NaN result (or a
// NaN centre) is the signal to redo that cell with the general axis_grad - map borders and nodata only.
__device__ __forceinline__ float4 load_quad(const float* __restrict__ h, int x, int y0, int nx, int ny) {
  if (x < 0 || x >= nx) return make_float4(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F);
  return __ldg(reinterpret_cast<const float4*>(h + (size_t)x * ny + y0));
}

constexpr int kNormalAhead = 1;  // rows requested ahead of their use; 2, 3, 4 measured 4, 10, 12 % slower (registers)
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_normal_sweep(const float* __restrict__ h,
                                                                      float* __restrict__ out, int nx, int ny,
                                                                      int band_rows, float rx, float ry) {
  const int lane = threadIdx.x & 31;
  const int strips = (ny + 127) / 128;
  const long long wid = (long long)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  const int strip = (int)(wid % strips);
  const int x0 = (int)(wid / strips) * band_rows;
  const int y0 = strip * 128 + lane * 4;
  if (x0 >= nx || y0 >= ny) return;
  const int x1 = min(x0 + band_rows, nx);
  const bool okL = y0 >= 2, okR = y0 + 4 < ny;  // ny % 4 == 0: the right pair is inside the row or entirely out

  float4 rm2 = load_quad(h, x0 - 2, y0, nx, ny), rm1 = load_quad(h, x0 - 1, y0, nx, ny);
  float4 r0 = load_quad(h, x0, y0, nx, ny), rp1 = load_quad(h, x0 + 1, y0, nx, ny);
  // Everything an iteration consumes was requested kNormalAhead iterations earlier: row x+2 and the +-2 column taps
  // of row x.  Row indices past the band's lower halo or past the map are clamped to the last row the band reads
  // anyway (a cache hit); a row outside the map is then replaced by the NaN markers when it is first used.
  const int xlast = min(x1 + 1, nx - 1);
  auto quad_at = [&](int x) { return __ldg(reinterpret_cast<const float4*>(h + (size_t)min(x, xlast) * ny + y0)); };
  const float2 nan2 = make_float2(CUDART_NAN_F, CUDART_NAN_F);
  auto taps_l = [&](int x) { return okL ? __ldg(reinterpret_cast<const float2*>(h + (size_t)min(x, x1 - 1) * ny + y0 - 2)) : nan2; };
  auto taps_r = [&](int x) { return okR ? __ldg(reinterpret_cast<const float2*>(h + (size_t)min(x, x1 - 1) * ny + y0 + 4)) : nan2; };
  float4 qr[kNormalAhead];
  float2 ql[kNormalAhead], qh[kNormalAhead];
#pragma unroll
  for (int d = 0; d < kNormalAhead; ++d) {
    qr[d] = quad_at(x0 + 2 + d);
    ql[d] = taps_l(x0 + d);
    qh[d] = taps_r(x0 + d);
  }
#pragma unroll 5 + kNormalAhead
  for (int x = x0; x < x1; ++x) {
    const float4 ahead = quad_at(x + 2 + kNormalAhead);
    const float2 hl_ahead = taps_l(x + kNormalAhead), hr_ahead = taps_r(x + kNormalAhead);
    float4 rp2 = qr[0];
    const float2 hl = ql[0], hr = qh[0];
#pragma unroll
    for (int d = 0; d + 1 < kNormalAhead; ++d) { qr[d] = qr[d + 1]; ql[d] = ql[d + 1]; qh[d] = qh[d + 1]; }
    qr[kNormalAhead - 1] = ahead; ql[kNormalAhead - 1] = hl_ahead; qh[kNormalAhead - 1] = hr_ahead;
    if (x + 2 >= nx) rp2 = make_float4(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F);
    const float w[8] = {hl.x, hl.y, r0.x, r0.y, r0.z, r0.w, hr.x, hr.y};
    const float am2[4] = {rm2.x, rm2.y, rm2.z, rm2.w}, am1[4] = {rm1.x, rm1.y, rm1.z, rm1.w};
    const float ap1[4] = {rp1.x, rp1.y, rp1.z, rp1.w}, ap2[4] = {rp2.x, rp2.y, rp2.z, rp2.w};
    float o[12];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float f0 = w[j + 2];
      float dx = ((am2[j] - 8.0f * am1[j]) + (8.0f * ap1[j] - ap2[j])) * kTwelfth;
      float dy = ((w[j] - 8.0f * w[j + 1]) + (8.0f * w[j + 3] - w[j + 4])) * kTwelfth;
      if (dx != dx || dy != dy || f0 != f0) {
        dx = axis_grad(am2[j], am1[j], f0, ap1[j], ap2[j]);
        dy = axis_grad(w[j], w[j + 1], f0, w[j + 3], w[j + 4]);
      }
      unit_normal(dx, dy, rx, ry, o[3 * j], o[3 * j + 1], o[3 * j + 2]);
    }
    float4* q = reinterpret_cast<float4*>(out + 3 * ((size_t)x * ny + y0));
    __stcs(q, make_float4(o[0], o[1], o[2], o[3]));
    __stcs(q + 1, make_float4(o[4], o[5], o[6], o[7]));
    __stcs(q + 2, make_float4(o[8], o[9], o[10], o[11]));
    rm2 = rm1; rm1 = r0; r0 = rp1; rp1 = rp2;
  }
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

const char* soilb200_version(void) { return "soillib_b200 0.1 (sm_100a)"; }

int soilb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int soilb200_rng_state_bytes(void) { return (int)sizeof(curandStateXORWOW_t); }

uint64_t soilb200_launch_count(void) { return g_launches; }

void soilb200_param_default(soilb200_param_t* p) {  // erosion.hpp:20-56
  if (!p) return;
  p->maxage = 512; p->lrate = 1.0f; p->timeStep = 250.0f;
  p->exitSlope = 0.02f; p->uplift = 0.001f; p->rainfall = 1.0f; p->gravity = 9.81f; p->evapRate = 0.0002f;
  p->frictionFactor = 0.06f; p->fluvialExponent = 2.0f;
  p->suspensionRateFluvial = 4.5E-8f; p->depositionRateFluvial = 0.04f;
  p->suspensionRateDebris = 0.001f; p->depositionRateDebris = 0.01f; p->landslideRateDebris = 0.003f;
  p->critSlopeBedrock = 0.57f; p->critSlopeSediment = 0.3f; p->yieldStress = 0.001f;
  p->viscosityWater = 1E-6f; p->bedShearWater = 0.0075f; p->densityWater = 1.0f;
  p->viscosityDebris = 0.0f; p->bedShearDebris = 0.99f; p->densityDebris = 2.0f;
  p->force[0] = 0.0f; p->force[1] = 0.0f;
}

int soilb200_direction_f32(const float* height, int32_t* out, int nx, int ny, int edge, void* stream) {
  return launch_steepest<true>(height, out, nx, ny, edge, (cudaStream_t)stream);
}

int soilb200_steepest_f32(const float* height, int32_t* out, int nx, int ny, int edge, void* stream) {
  return launch_steepest<false>(height, out, nx, ny, edge, (cudaStream_t)stream);
}

int soilb200_seed_xorwow(void* states, int64_t n, uint64_t seed, uint64_t first_subsequence, uint64_t offset,
                         void* stream) {
  if (!states || n <= 0) return SOILB200_E_ARG;
  k_seed<<<blocks_for(n, 256), 256, 0, (cudaStream_t)stream>>>((curandStateXORWOW_t*)states, n, seed,
                                                               first_subsequence, offset);
  count_launch();
  return check_launch();
}

size_t soilb200_random_weighted_ws(int nx, int ny) {
  if (nx <= 0 || ny <= 0) return 0;
  return sizeof(rw_header) + 6 * rw_pitch((long long)nx * ny) * sizeof(unsigned);
}

int soilb200_random_weighted_f32(const float* height, int32_t* out, int nx, int ny, int edge, uint64_t seed,
                                 uint64_t offset, float temperature, void* ws, size_t ws_bytes, void* stream) {
  return soilb200_random_weighted_strip_f32(height, out, nx, ny, edge, seed, offset, temperature, 0, ws, ws_bytes, stream);
}

int soilb200_random_weighted_strip_f32(const float* height, int32_t* out, int nx, int ny, int edge, uint64_t seed,
                                       uint64_t offset, float temperature, uint64_t first_cell, void* ws,
                                       size_t ws_bytes, void* stream) {
  if (!height || !out || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if (edge != SOILB200_D4 && edge != SOILB200_D8) return SOILB200_E_EDGE;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  if (ws && ws_bytes < soilb200_random_weighted_ws(nx, ny)) return SOILB200_E_WORKSPACE;
  cudaStream_t s = (cudaStream_t)stream;
  const long long N = (long long)nx * ny;
  rw_args a;
  a.hdr = (const rw_header*)ws;
  a.planes = ws ? (unsigned*)((char*)ws + sizeof(rw_header)) : nullptr;
  a.pitch = rw_pitch(N);
  a.seed = seed; a.offset = offset; a.T = temperature; a.base = first_cell;
  if (ws) {  // make the stored streams stand at (seed, offset): a no-op when the previous call left them there
    const long long threads = (N + 16LL * kSeedChain - 1) / (16LL * kSeedChain) * 16;
    k_rw_seed<<<blocks_for(threads, 128), 128, 0, s>>>(a.hdr, a.planes, a.pitch, N, seed, offset, first_cell, nx, ny);
    k_rw_header<<<1, 1, 0, s>>>((rw_header*)ws, seed, offset, first_cell, nx, ny);
    count_launch(2);
  }
  if (vector_ok(height, out, ny)) {
    static const int resident = resident_blocks(k_sweep_gibbs<8>, kWarpsPerBlock * 32);
    const int br = sweep_band_rows(nx, ny, resident, 2);
    if (edge == SOILB200_D4) k_sweep_gibbs<4><<<sweep_blocks(nx, ny, br), kWarpsPerBlock * 32, 0, s>>>(height, out, nx, ny, br, a);
    else k_sweep_gibbs<8><<<sweep_blocks(nx, ny, br), kWarpsPerBlock * 32, 0, s>>>(height, out, nx, ny, br, a);
  } else {
    if (edge == SOILB200_D4) k_scalar_gibbs<4><<<blocks_for(N, 256), 256, 0, s>>>(height, out, nx, ny, a);
    else k_scalar_gibbs<8><<<blocks_for(N, 256), 256, 0, s>>>(height, out, nx, ny, a);
  }
  count_launch();
  if (ws) {
    k_rw_header<<<1, 1, 0, s>>>((rw_header*)ws, seed, offset + 1, first_cell, nx, ny);
    count_launch();
  }
  return check_launch();
}

int soilb200_selftest_div_sqrt2(uint64_t* mismatches_device, void* stream) {
  if (!mismatches_device) return SOILB200_E_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(mismatches_device, 0, sizeof(uint64_t), s);
  if (e != cudaSuccess) return (int)e;
  k_selftest_div_sqrt2<<<sm_count() * 16, 256, 0, s>>>((unsigned long long*)mismatches_device);
  count_launch();
  return check_launch();
}

int soilb200_normal_f32(const float* height, float* out, int nx, int ny, const float scale[3], void* stream) {
  if (!height || !out || !scale || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  const float rx = scale[2] / scale[0], ry = scale[2] / scale[1];
  if (vector_ok(height, out, ny)) {
    static const int resident = resident_blocks(k_normal_sweep, kWarpsPerBlock * 32);
    const int br = sweep_band_rows(nx, ny, resident, 4);
    k_normal_sweep<<<sweep_blocks(nx, ny, br), kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(height, out, nx, ny, br, rx, ry);
  } else {
    const dim3 grid((ny + 31) / 32, (nx + 7) / 8);
    k_normal<<<grid, 256, 0, (cudaStream_t)stream>>>(height, out, nx, ny, rx, ry);
  }
  count_launch();
  return check_launch();
}

}  // extern "C"
