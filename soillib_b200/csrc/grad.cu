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

// one axis of grad.cu:58-69: central difference, else backward, else forward, else 0
__device__ __forceinline__ float grad_axis(float c, float m, float p, float s) {
  float g = 0.5f * (p - m) / s;
  if (g != g) {  // map border or nodata: the one-sided candidates, computed only here
    g = (c - m) / s;
    if (g != g) g = (p - c) / s;
    if (g != g) g = 0.0f;
  }
  return g;
}

// grad.cu:119-129 per axis.  Dividing by a positive scale is monotone under round-to-nearest, so
// max(a / s, b / s) == max(a, b) / s bit for bit: one division per axis (a negative scale flips the order).
__device__ __forceinline__ float negslope_axis(float c, float m, float p, float s) {
  const float a = c - m, b = c - p;                       // NaN where the neighbour does not exist (or is nodata)
  const float q = (s > 0.0f ? fmaxf(a, b) : fminf(a, b)) / s;  // fmaxf / fminf drop a NaN operand
  return 0.0f < q ? q : 0.0f;                             // glm::max(0, q); NaN -> 0
}

// grad.cu:22-89
__global__ void __launch_bounds__(256) k_gradient(const float* __restrict__ h, float2* __restrict__ out, int nx, int ny,
                                                  float sx, float sy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
  const h5 w = load5(h, c, nx, ny);
  out[c.n] = make_float2(grad_axis(w.c, w.n0, w.p0, sx), grad_axis(w.c, w.zn, w.zp, sy));
}

// grad.cu:102-130
__global__ void __launch_bounds__(256) k_negslope(const float* __restrict__ h, float* __restrict__ out, int nx, int ny,
                                                  float sx, float sy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
  const h5 w = load5(h, c, nx, ny);
  out[c.n] = len2(negslope_axis(w.c, w.n0, w.p0, sx), negslope_axis(w.c, w.zn, w.zp, sy));
}

// grad.cu:137-176, D interleaved components per cell
template <int D>
__global__ void __launch_bounds__(256) k_laplacian(const float* __restrict__ in, float* __restrict__ out, int nx, int ny,
                                                   float hx, float hy) {
  const cell c = my_cell(nx, ny);
  if (!c.in) return;
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

// ---- the same three stencils as register-window sweeps (flow.cu's shape of the work) ------------------------
// A warp sweeps a 128-column strip down a band of rows; a lane owns 4 consecutive cells (one 16-byte load per
// row, the two halo cells are L1 hits on the neighbouring lanes' lines), the rows above and below stay in
// registers, results leave as 16-byte stores.  Used when the row length is a multiple of 4 and the bases are
// 16-byte aligned; the one-thread-per-cell kernels above cover the rest.
constexpr int kSweepWarps = 4;
constexpr int kSweepBand = 32;

struct row6 {
  float w[6];  // columns y0-1 .. y0+4; NaN where the map ends
};
__device__ __forceinline__ row6 load_row6(const float* __restrict__ h, int x, int y0, int nx, int ny, bool okL,
                                          bool okR) {
  row6 r;
#pragma unroll
  for (int i = 0; i < 6; ++i) r.w[i] = CUDART_NAN_F;
  if (x >= 0 && x < nx) {
    const float* p = h + (size_t)x * ny + y0;
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    r.w[1] = v.x; r.w[2] = v.y; r.w[3] = v.z; r.w[4] = v.w;
    if (okL) r.w[0] = __ldg(p - 1);
    if (okR) r.w[5] = __ldg(p + 4);
  }
  return r;
}

struct sweep_pos {
  int x0, x1, y0;
  bool in, okL, okR;
};
__device__ __forceinline__ sweep_pos my_band(int nx, int ny, int band_rows) {
  sweep_pos p;
  const int strips = (ny + 127) / 128;
  const long long wid = (long long)blockIdx.x * kSweepWarps + (threadIdx.x >> 5);
  p.x0 = (int)(wid / strips) * band_rows;
  p.y0 = (int)(wid % strips) * 128 + (threadIdx.x & 31) * 4;
  p.in = p.x0 < nx && p.y0 < ny;
  p.x1 = min(p.x0 + band_rows, nx);
  p.okL = p.y0 > 0;
  p.okR = p.y0 + 4 < ny;
  return p;
}

// Row feeder of the sweeps: the rows below the window travel through a private shared-memory ring per warp,
// kFeedSlots slots of one 128-column row (+ one halo cell each side) filled by cp.async copies (each lane its own
// 4 cells, lanes 0 / 31 the halo cell) kFeedSlots - 1 rows ahead of their use - the rows in flight that the HBM
// latency asks for, without holding registers (flow.cu's k_steepest measured it: long_scoreboard 5.5 -> 0.6 cycles
// per issue; here: gradient 0.175 -> 0.166 ms, negslope 0.146 -> 0.142 ms at 8192^2).  No block-level barrier: a warp waits on its own copy group, __syncwarp()s, reads; a slot is refilled
// one iteration after it was read, when every lane is past that read.
constexpr int kFeedSlots = 8;  // a power of two
constexpr int kFeedRow = 4 + 128 + 4;  // [.. left halo | 128 columns | right halo ..] floats, 16-byte granules
struct row_feed {
  unsigned base;      // shared-memory address of this lane's 4 cells in slot 0
  const float* src0;  // first owned cell of row 0
  int ny, xlast;      // refills past the band's lower halo (or the map) re-read the last row the band needs
  unsigned lanes;
  bool okL, okR, haloL, haloR;
  __device__ __forceinline__ void issue(int x, unsigned slot) const {
    const float* src = src0 + min(x, xlast) * ny;  // nx * ny < 2^31 is checked by the launchers
    const unsigned dst = base + slot * (unsigned)(kFeedRow * sizeof(float));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
    if (haloL) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst - 4u), "l"(src - 1) : "memory");
    if (haloR) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + 16u), "l"(src + 4) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // rows first .. first + kFeedSlots - 2 into slots 0 ..
  __device__ __forceinline__ void prime(float (*ring)[kFeedRow], const float* h, const sweep_pos& p, int nx, int ny_,
                                        unsigned lanes_, int first) {
    const int lane = threadIdx.x & 31;
    base = (unsigned)__cvta_generic_to_shared(&ring[0][4 + 4 * lane]);
    src0 = h + p.y0;
    ny = ny_;
    xlast = min(p.x1, nx - 1);
    lanes = lanes_;
    okL = p.okL; okR = p.okR;
    haloL = lane == 0 && p.okL; haloR = lane == 31 && p.okR;
#pragma unroll
    for (int d = 0; d < kFeedSlots - 1; ++d) issue(first + d, (unsigned)d);
  }
  // the i-th row since prime(), which is row x of the map (NaN markers when x is outside it)
  __device__ __forceinline__ row6 take(unsigned i, int x, int nx) const {
    asm volatile("cp.async.wait_group %0;" ::"n"(kFeedSlots - 2) : "memory");
    __syncwarp(lanes);
    row6 r;
    const unsigned a = base + (i & (kFeedSlots - 1)) * (unsigned)(kFeedRow * sizeof(float));
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.w[1]), "=f"(r.w[2]), "=f"(r.w[3]), "=f"(r.w[4]) : "r"(a) : "memory");
    r.w[0] = r.w[5] = CUDART_NAN_F;
    if (okL) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r.w[0]) : "r"(a - 4u) : "memory");
    if (okR) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r.w[5]) : "r"(a + 16u) : "memory");
    issue(x + kFeedSlots - 1, (i + kFeedSlots - 1) & (kFeedSlots - 1));
    if (x >= nx) {
#pragma unroll
      for (int k = 0; k < 6; ++k) r.w[k] = CUDART_NAN_F;
    }
    return r;
  }
  __device__ __forceinline__ void drain() const { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
};


__global__ void __launch_bounds__(kSweepWarps * 32) k_gradient_sweep(const float* __restrict__ h,
                                                                     float* __restrict__ out, int nx, int ny,
                                                                     int band_rows, float sx, float sy) {
  const sweep_pos p = my_band(nx, ny, band_rows);
  const unsigned lanes = __ballot_sync(0xffffffffu, p.in);  // the lanes that sweep; the others leave
  if (!p.in) return;
  __shared__ __align__(16) float ring[kSweepWarps][kFeedSlots][kFeedRow];
  row_feed feed;
  feed.prime(ring[threadIdx.x >> 5], h, p, nx, ny, lanes, p.x0 + 1);
  unsigned fed = 0;
  row6 up = load_row6(h, p.x0 - 1, p.y0, nx, ny, p.okL, p.okR);
  row6 cu = load_row6(h, p.x0, p.y0, nx, ny, p.okL, p.okR);
#pragma unroll 2
  for (int x = p.x0; x < p.x1; ++x) {
    const row6 dn = feed.take(fed++, x + 1, nx);
    float g[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      g[2 * j] = grad_axis(cu.w[j + 1], up.w[j + 1], dn.w[j + 1], sx);
      g[2 * j + 1] = grad_axis(cu.w[j + 1], cu.w[j], cu.w[j + 2], sy);
    }
    float4* q = reinterpret_cast<float4*>(out + 2 * ((size_t)x * ny + p.y0));
    __stcs(q, make_float4(g[0], g[1], g[2], g[3]));
    __stcs(q + 1, make_float4(g[4], g[5], g[6], g[7]));
    up = cu;
    cu = dn;
  }
  feed.drain();  // nothing may land in the ring after the block is gone
}

__global__ void __launch_bounds__(kSweepWarps * 32) k_negslope_sweep(const float* __restrict__ h,
                                                                     float* __restrict__ out, int nx, int ny,
                                                                     int band_rows, float sx, float sy) {
  const sweep_pos p = my_band(nx, ny, band_rows);
  const unsigned lanes = __ballot_sync(0xffffffffu, p.in);  // the lanes that sweep; the others leave
  if (!p.in) return;
  __shared__ __align__(16) float ring[kSweepWarps][kFeedSlots][kFeedRow];
  row_feed feed;
  feed.prime(ring[threadIdx.x >> 5], h, p, nx, ny, lanes, p.x0 + 1);
  unsigned fed = 0;
  row6 up = load_row6(h, p.x0 - 1, p.y0, nx, ny, p.okL, p.okR);
  row6 cu = load_row6(h, p.x0, p.y0, nx, ny, p.okL, p.okR);
#pragma unroll 2
  for (int x = p.x0; x < p.x1; ++x) {
    const row6 dn = feed.take(fed++, x + 1, nx);
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      r[j] = len2(negslope_axis(cu.w[j + 1], up.w[j + 1], dn.w[j + 1], sx),
                  negslope_axis(cu.w[j + 1], cu.w[j], cu.w[j + 2], sy));
    __stcs(reinterpret_cast<float4*>(out + (size_t)x * ny + p.y0), make_float4(r[0], r[1], r[2], r[3]));
    up = cu;
    cu = dn;
  }
  feed.drain();  // nothing may land in the ring after the block is gone
}

// grad.cu:137-176 for one component: neighbours outside the map continue the centre value (by POSITION, a NaN
// inside the map is data), hx = 1/sx/sx and hy = 1/sy/sy rounded as the reference rounds them
__global__ void __launch_bounds__(kSweepWarps * 32) k_laplacian_sweep(const float* __restrict__ in,
                                                                      float* __restrict__ out, int nx, int ny,
                                                                      int band_rows, float hx, float hy) {
  const sweep_pos p = my_band(nx, ny, band_rows);
  if (!p.in) return;
  row6 up = load_row6(in, p.x0 - 1, p.y0, nx, ny, p.okL, p.okR);
  row6 cu = load_row6(in, p.x0, p.y0, nx, ny, p.okL, p.okR);
#pragma unroll 2
  for (int x = p.x0; x < p.x1; ++x) {
    const row6 dn = load_row6(in, x + 1, p.y0, nx, ny, p.okL, p.okR);  // the ring feeder measured 1 % slower here
    const bool xm = x > 0, xp = x + 1 < nx;
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool ym = j > 0 || p.okL, yp = j < 3 || p.okR;
      const float v00 = cu.w[j + 1];
      const float vn0 = xm ? up.w[j + 1] : v00, vp0 = xp ? dn.w[j + 1] : v00;
      const float v0n = ym ? cu.w[j] : v00, v0p = yp ? cu.w[j + 2] : v00;
      const float vnn = (xm && ym) ? up.w[j] : v00, vpp = (xp && yp) ? dn.w[j + 2] : v00;
      const float vpn = (xp && ym) ? dn.w[j] : v00, vnp = (xm && yp) ? up.w[j + 2] : v00;
      const float LH = (vn0 - v00) * hx + (vp0 - v00) * hx + (v0n - v00) * hy + (v0p - v00) * hy;
      const float LD = 0.5f * (vnn - v00) * hx + 0.5f * (vpp - v00) * hx + 0.5f * (vpn - v00) * hy + 0.5f * (vnp - v00) * hy;
      r[j] = 0.5f * LH + 0.5f * LD;
    }
    __stcs(reinterpret_cast<float4*>(out + (size_t)x * ny + p.y0), make_float4(r[0], r[1], r[2], r[3]));
    up = cu;
    cu = dn;
  }
}

// graph.cu:270-296, 4 cells per thread (16-byte loads of the graph and the heights, one 16-byte store).  The row of
// the receiver is almost always x-1, x or x+1: those are tested before falling back to the integer division.
__device__ __forceinline__ float slope_cell(const float* __restrict__ h, int n, int x, int y, int next, float hc, int nx,
                                            int ny, float sx, float sy) {
  if (next < 0 || next == n || next >= nx * ny) return 0.0f;
  const int r0 = (x - 1) * ny;
  int nr;
  if (next >= r0 && next < r0 + 3 * ny) nr = x - 1 + (next >= r0 + ny) + (next >= r0 + 2 * ny);
  else nr = next / ny;
  const int nc = next - nr * ny;
  const float dx = sx * ((float)nr - (float)x), dy = sy * ((float)nc - (float)y);
  return (__ldg(h + next) - hc) / len2(dx, dy);
}
__global__ void __launch_bounds__(256) k_slope_quad(const float* __restrict__ h, const int32_t* __restrict__ flow,
                                                    float* __restrict__ out, int nx, int ny, float sx, float sy) {
  const int quads = ny >> 2;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)nx * quads) return;
  const int x = (int)(t / quads), y0 = (int)(t - (long long)x * quads) * 4;
  const int n0 = x * ny + y0;
  const int4 f = __ldcs(reinterpret_cast<const int4*>(flow + n0));
  const float4 c = __ldg(reinterpret_cast<const float4*>(h + n0));
  float4 r;
  r.x = slope_cell(h, n0, x, y0, f.x, c.x, nx, ny, sx, sy);
  r.y = slope_cell(h, n0 + 1, x, y0 + 1, f.y, c.y, nx, ny, sx, sy);
  r.z = slope_cell(h, n0 + 2, x, y0 + 2, f.z, c.z, nx, ny, sx, sy);
  r.w = slope_cell(h, n0 + 3, x, y0 + 3, f.w, c.w, nx, ny, sx, sy);
  __stcs(reinterpret_cast<float4*>(out + n0), r);
}

inline bool sweep_ok(const void* a, const void* b, int ny) {
  return (ny % 4 == 0) && ((reinterpret_cast<uintptr_t>(a) & 15) == 0) && ((reinterpret_cast<uintptr_t>(b) & 15) == 0);
}
inline int sweep_rows(int nx, int ny) {  // fewer rows per warp on small maps so that the grid still fills the GPU
  const long long strips = (ny + 127) / 128;
  long long rows = (long long)nx * strips / ((long long)sm_count() * 48);
  return (int)(rows > kSweepBand ? kSweepBand : (rows < 4 ? 4 : rows));
}
inline unsigned sweep_grid(int nx, int ny, int rows) {
  const long long strips = (ny + 127) / 128, bands = (nx + rows - 1) / rows;
  return (unsigned)((strips * bands + kSweepWarps - 1) / kSweepWarps);
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

// ---- the blur for one component, register-blocked ----------------------------------------------------------
// The arithmetic floor of a pass without multiply-add contraction is 33 multiplies + 33 adds per cell; k_blur adds
// 33 loads and the index clamping to that.  Here the 33 weights and a sliding window of inputs live in registers:
//  - along dim 1 a lane owns 4 consecutive outputs and reads their 36 inputs as nine 16-byte loads;
//  - along dim 0 a thread owns one column and keeps the last 33 rows in a register ring (the loop is unrolled over
//    the 33 phases of the ring so that every index is static), one coalesced 4-byte load and store per row.
// Same sum order as the reference (taps k = -16..16 ascending), so the results are those of k_blur bit for bit.
constexpr int kTaps = 2 * kBlurWindow + 1;

__device__ __forceinline__ void blur_weights(float (&w)[kTaps], float* smem, float sigma) {
  if (threadIdx.x < kTaps) {
    const int k = (int)threadIdx.x - kBlurWindow;
    const float Z = sqrtf(2.0f * 3.14159265f) * sigma;
    smem[threadIdx.x] = __expf(-0.5f * (k / sigma) * (k / sigma)) / Z;  // filter.cu:45-46
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kTaps; ++i) w[i] = smem[i];
}

constexpr int kBlurYRows = 8;  // rows per warp in the dim-1 pass
__global__ void __launch_bounds__(128) k_blur_y(const float* __restrict__ src, float* __restrict__ dst, int nx, int ny,
                                                float sigma) {
  __shared__ float wsm[kTaps];
  float w[kTaps];
  blur_weights(w, wsm, sigma);
  const int strips = (ny + 127) / 128;
  const long long wid = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int x0 = (int)(wid / strips) * kBlurYRows;
  const int y0 = (int)(wid % strips) * 128 + (threadIdx.x & 31) * 4;
  if (x0 >= nx || y0 >= ny) return;
  const int x1 = min(x0 + kBlurYRows, nx);
  const bool inner = y0 >= kBlurWindow && y0 + 4 + kBlurWindow <= ny;
  for (int x = x0; x < x1; ++x) {
    const float* row = src + (size_t)x * ny;
    float win[4 + 2 * kBlurWindow];  // inputs y0-16 .. y0+19
    if (inner) {
#pragma unroll
      for (int i = 0; i < (4 + 2 * kBlurWindow) / 4; ++i) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(row + y0 - kBlurWindow) + i);
        win[4 * i] = v.x; win[4 * i + 1] = v.y; win[4 * i + 2] = v.z; win[4 * i + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4 + 2 * kBlurWindow; ++i) win[i] = __ldg(row + min(max(y0 - kBlurWindow + i, 0), ny - 1));
    }
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float val = 0.0f;
#pragma unroll
      for (int k = 0; k < kTaps; ++k) val += win[j + k] * w[k];
      r[j] = val;
    }
    *reinterpret_cast<float4*>(dst + (size_t)x * ny + y0) = make_float4(r[0], r[1], r[2], r[3]);
  }
}

constexpr int kBlurXBand = 8 * kTaps;  // output rows per thread in the dim-0 pass on large maps (32 extra input rows per band)
__global__ void __launch_bounds__(128) k_blur_x(const float* __restrict__ src, float* __restrict__ dst, int nx, int ny,
                                                int band, float sigma) {
  __shared__ float wsm[kTaps];
  float w[kTaps];
  blur_weights(w, wsm, sigma);
  const int y = blockIdx.x * 128 + threadIdx.x;
  const int x0 = blockIdx.y * band;
  if (y >= ny || x0 >= nx) return;
  const int x1 = min(x0 + band, nx);
  auto in = [&](int x) { return __ldg(src + (size_t)min(max(x, 0), nx - 1) * ny + y); };
  // ring slot of input row x0 + i is (i + 16) % 33, i = -16 .. ; output row x0 + p reads slots p .. p + 32 (mod 33)
  float ring[kTaps];
#pragma unroll
  for (int i = 0; i < kTaps - 1; ++i) ring[i] = in(x0 - kBlurWindow + i);  // rows x0-16 .. x0+15
  for (int xb = x0; xb < x1; xb += kTaps) {
#pragma unroll
    for (int p = 0; p < kTaps; ++p) {
      const int x = xb + p;
      if (x < x1) {
        ring[(p + kTaps - 1) % kTaps] = in(x + kBlurWindow);  // the newest row replaces row x - 17
        float val = 0.0f;
#pragma unroll
        for (int k = 0; k < kTaps; ++k) val += ring[(p + k) % kTaps] * w[k];
        dst[(size_t)x * ny + y] = val;
      }
    }
  }
}

inline bool bad2(const float* s) { return !s || !(s[0] == s[0]) || !(s[1] == s[1]); }

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

int soilb200_gradient_f32(const float* height, float* out, int nx, int ny, const float scale[2], void* stream) {
  if (!height || !out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  if (sweep_ok(height, out, ny)) {
    const int rows = sweep_rows(nx, ny);
    k_gradient_sweep<<<sweep_grid(nx, ny, rows), kSweepWarps * 32, 0, s>>>(height, out, nx, ny, rows, scale[0], scale[1]);
  } else {
    k_gradient<<<cell_grid(nx, ny), 256, 0, s>>>(height, (float2*)out, nx, ny, scale[0], scale[1]);
  }
  count_launch();
  return check_launch();
}

int soilb200_negslope_f32(const float* height, float* out, int nx, int ny, const float scale[2], void* stream) {
  if (!height || !out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  if (sweep_ok(height, out, ny)) {
    const int rows = sweep_rows(nx, ny);
    k_negslope_sweep<<<sweep_grid(nx, ny, rows), kSweepWarps * 32, 0, s>>>(height, out, nx, ny, rows, scale[0], scale[1]);
  } else {
    k_negslope<<<cell_grid(nx, ny), 256, 0, s>>>(height, out, nx, ny, scale[0], scale[1]);
  }
  count_launch();
  return check_launch();
}

int soilb200_laplacian_f32(const float* in, float* out, int nx, int ny, int channels, const float scale[2],
                           void* stream) {
  if (!in || !out || in == out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if (channels != 1 && channels != 2) return SOILB200_E_ARG;  // the reference instantiates D = 1, 2 (grad.cu:193-199)
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  const float hx = (1.0f / scale[0] / scale[0]), hy = (1.0f / scale[1] / scale[1]);  // grad.cu:163-164, rounded once
  if (channels == 1 && sweep_ok(in, out, ny)) {
    const int rows = sweep_rows(nx, ny);
    k_laplacian_sweep<<<sweep_grid(nx, ny, rows), kSweepWarps * 32, 0, s>>>(in, out, nx, ny, rows, hx, hy);
  } else if (channels == 1) k_laplacian<1><<<cell_grid(nx, ny), 256, 0, s>>>(in, out, nx, ny, hx, hy);
  else k_laplacian<2><<<cell_grid(nx, ny), 256, 0, s>>>(in, out, nx, ny, hx, hy);
  count_launch();
  return check_launch();
}

int soilb200_slope_f32(const float* height, const int32_t* flow, float* out, int nx, int ny, const float scale[2],
                       void* stream) {
  if (!height || !flow || !out || nx <= 0 || ny <= 0 || bad2(scale)) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  if (sweep_ok(height, out, ny) && (reinterpret_cast<uintptr_t>(flow) & 15) == 0) {
    k_slope_quad<<<blocks_for((long long)nx * (ny >> 2), 256), 256, 0, s>>>(height, flow, out, nx, ny, scale[0], scale[1]);
  } else {
    k_slope<<<cell_grid(nx, ny), 256, 0, s>>>(height, flow, out, nx, ny, scale[0], scale[1]);
  }
  count_launch();
  return check_launch();
}

int soilb200_gaussian_blur_f32(float* inout, float* tmp, int nx, int ny, int channels, float sigma, void* stream) {
  if (!inout || !tmp || inout == tmp || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if (channels != 1 && channels != 2) return SOILB200_E_ARG;  // filter.cu:79-88
  if ((long long)nx * ny * channels > 0x7fffffffLL) return SOILB200_E_LIMIT;
  cudaStream_t s = (cudaStream_t)stream;
  const dim3 grid((ny * channels + 31) / 32, (nx + 7) / 8);
  // rows per thread of the dim-0 pass: 264 on large maps, fewer (a multiple of the ring length) when that would
  // leave SMs without a block
  int band = kBlurXBand;
  {
    const int col_blocks = (ny + 127) / 128;
    const int want_bands = (2 * sm_count() + col_blocks - 1) / col_blocks;
    const int rows = (nx + want_bands - 1) / want_bands;
    band = ((rows + kTaps - 1) / kTaps) * kTaps;
    band = band > kBlurXBand ? kBlurXBand : band;
  }
  if (channels == 1 && (nx + band - 1) / band <= 65535) {
    k_blur_x<<<dim3((ny + 127) / 128, (nx + band - 1) / band), 128, 0, s>>>(inout, tmp, nx, ny, band, sigma);
  } else {
    k_blur<<<grid, 256, 0, s>>>(inout, tmp, nx, ny, channels, sigma, true);
  }
  if (channels == 1 && sweep_ok(inout, tmp, ny)) {
    const long long warps = (long long)((ny + 127) / 128) * ((nx + kBlurYRows - 1) / kBlurYRows);
    k_blur_y<<<(unsigned)((warps + 3) / 4), 128, 0, s>>>(tmp, inout, nx, ny, sigma);
  } else {
    k_blur<<<grid, 256, 0, s>>>(tmp, inout, nx, ny, channels, sigma, false);
  }
  count_launch(2);
  return check_launch();
}

}  // extern "C"
