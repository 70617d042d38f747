// accumulate_tiled.cu - upstream accumulation (no decay) resolved tile by tile in shared memory.
//
// Reference: source/soillib/model/graph/graph.cu:321-593 (soil::accumulate).  Same sums as
// accumulate.cu's global walk, but the long serial chains (a river is tens of thousands of cells) are
// walked at shared-memory latency inside 64x64 tiles, and only tile-boundary cells meet in global memory:
//
//   A  per tile (smem): in-tile donor counts, then leaf-started walkers give every cell the sum of its
//      IN-TILE upstream area.  Cells whose receiver lies in another tile are "exit" cells; their local
//      totals go to `out`; every perimeter cell records which exit cell (if any) it drains to.
//   B  reduced forest over exit cells only (~T/2 x fewer hops than cells): exit e -> the exit cell that
//      e's receiver drains to.  Same counter walk as accumulate.cu, totals added with float atomics.
//   C  per tile again: perimeter cells start from source + the finished totals of the exit cells that
//      flow into them from outside, walkers redo the tile, every non-exit cell is written.
//
// HBM traffic: A reads graph + source (8 B/cell), C reads them again and writes out (12 B/cell); B touches
// perimeter cells only.  Scratch 12 B/cell (root, pending, next).  Sums of integer-valued sources are
// exact; real-valued sources differ from the global walk in summation order only (phase B uses atomics).
#include "common.cuh"

namespace sb {
namespace {

constexpr int T = 64;               // tile edge
constexpr int TT = T * T;
#ifndef SOILB200_TILE_THREADS
#define SOILB200_TILE_THREADS 512
#endif
#ifndef SOILB200_WALK_STEPS
#define SOILB200_WALK_STEPS 16
#endif
constexpr int kThreads = SOILB200_TILE_THREADS;
constexpr int kCellsPerThread = TT / kThreads;  // 16
constexpr unsigned char kNone = 0xFF;

// neighbour code k of the edge n -> r, or -1 if r is not an in-map K-neighbour of n (graph.cu:336-346)
// graph.hpp:35-44 without the constant-bank tables (an indexed constant load serialises over the distinct k of a
// warp): the x and y components of shift_k, two bits each, biased by one
__device__ __forceinline__ int shift_dx(int k) { return (int)((0xA094u >> (2 * k)) & 3u) - 1; }
__device__ __forceinline__ int shift_dy(int k) { return (int)((0x8861u >> (2 * k)) & 3u) - 1; }

template <int K>
__device__ __forceinline__ int edge_code(int r, int n, int y, int nx, int ny) {
  if (r < 0 || r >= nx * ny) return -1;
  const int d = r - n;
  int k = -1;
  if (d == -ny) k = 0;
  else if (d == -1) k = 1;
  else if (d == 1) k = 2;
  else if (d == ny) k = 3;
  else if (K == 8) {
    if (d == -ny - 1) k = 4;
    else if (d == -ny + 1) k = 5;
    else if (d == ny - 1) k = 6;
    else if (d == ny + 1) k = 7;
  }
  if (k < 0) return -1;
  const int dy = shift_dy(k);
  if ((dy < 0 && y == 0) || (dy > 0 && y == ny - 1)) return -1;  // would wrap around a row end
  return k;
}

// shift_k as an offset inside the T x T tile (graph.hpp:35-44: -T, -1, 1, T, -T-1, -T+1, T-1, T+1), one signed byte
// per k packed in a 64-bit constant: two ALU instructions instead of an indexed constant-bank load, which
// serialises when the lanes of a warp ask for different k
__device__ __forceinline__ int tile_offset(int k) {
  static_assert(64 + 1 <= 127, "tile offsets must fit a signed byte");
  constexpr unsigned long long packed =
      ((unsigned long long)(unsigned char)(-64) << 0) | ((unsigned long long)(unsigned char)(-1) << 8) |
      ((unsigned long long)(unsigned char)(1) << 16) | ((unsigned long long)(unsigned char)(64) << 24) |
      ((unsigned long long)(unsigned char)(-65) << 32) | ((unsigned long long)(unsigned char)(-63) << 40) |
      ((unsigned long long)(unsigned char)(63) << 48) | ((unsigned long long)(unsigned char)(65) << 56);
  return (int)(signed char)(packed >> (8 * k));
}
// one shared-memory atomic per group of lanes that push together instead of one per lane
__device__ __forceinline__ int push_slot(int* count) {
  const unsigned m = __activemask();
  const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(count, __popc(m));
  base = __shfl_sync(m, base, leader);
  return base + __popc(m & ((1u << lane) - 1u));
}
constexpr int kWalkSteps = SOILB200_WALK_STEPS;  // cells one walker finishes before it re-queues the chain (warp load balance)

struct tile_smem {
  float val[TT];
  unsigned pend32[TT / 4];      // one byte per cell: donors still pending
  unsigned char code[TT];       // 0..7 receiver inside the tile, 8..15 receiver in another tile, 0xFF none
  unsigned char donors[TT];     // bit k set: the in-tile cell at c - shift[k] flows into c
  unsigned short queue[2][TT];  // ready cells of this round / of the next round
  int count[2];
};

// perimeter cell p of an h x w tile, each cell exactly once; returns false past the end
__device__ __forceinline__ bool perimeter_cell(int p, int h, int w, int& i, int& j) {
  if (p < w) { i = 0; j = p; return true; }
  p -= w;
  if (h > 1) {
    if (p < w) { i = h - 1; j = p; return true; }
    p -= w;
  }
  const int inner = h > 2 ? h - 2 : 0;
  if (p < inner) { i = 1 + p; j = 0; return true; }
  p -= inner;
  if (w > 1 && p < inner) { i = 1 + p; j = w - 1; return true; }
  return false;
}

// sum of finished totals of exit cells of OTHER tiles that flow into cell (x,y) = n
template <int K>
__device__ __forceinline__ float external_inflow(const int32_t* __restrict__ graph, const float* out, int x, int y,
                                                 int n, int x0, int y0, int h, int w, int nx, int ny) {
  float s = 0.0f;
#pragma unroll
  for (int k = 0; k < K; ++k) {  // slot order: the donor that reaches n by shift k sits at n - shift[k]
    const int qx = x - c_shift_x[k], qy = y - c_shift_y[k];
    if (qx < 0 || qy < 0 || qx >= nx || qy >= ny) continue;
    if (qx >= x0 && qx < x0 + h && qy >= y0 && qy < y0 + w) continue;  // in-tile donors come from the walk
    const int q = qx * ny + qy;
    if (edge_code<K>(__ldg(graph + q), q, qy, nx, ny) == k) s += __ldcg(out + q);
  }
  return s;
}

template <int K, bool FINAL>
__global__ void __launch_bounds__(kThreads) k_tile(const int32_t* __restrict__ graph, const float* __restrict__ src,
                                                   float* out, int* __restrict__ root, int* __restrict__ pend2,
                                                   int nx, int ny, int tiles_y) {
  __shared__ tile_smem sm;
  volatile float* val = sm.val;
  unsigned char* pend8 = reinterpret_cast<unsigned char*>(sm.pend32);
  const int x0 = (blockIdx.x / tiles_y) * T, y0 = (blockIdx.x % tiles_y) * T;
  const int h = min(T, nx - x0), w = min(T, ny - y0);
  const int tid = threadIdx.x;

  // ---- load: receiver codes and sources; every cell with an in-tile receiver sets its bit in the
  // receiver's donor mask (bit k: the cell at receiver - shift[k], i.e. the reference's donor slot k)
  unsigned* donors32 = reinterpret_cast<unsigned*>(sm.donors);
  for (int c4 = tid; c4 < TT / 4; c4 += kThreads) donors32[c4] = 0u;
  if (tid < 2) sm.count[tid] = 0;
  __syncthreads();
#pragma unroll 4
  for (int m = 0; m < kCellsPerThread; ++m) {
    const int c = tid + m * kThreads, i = c / T, j = c % T;
    unsigned char cd = kNone;
    float v = 0.0f;
    if (i < h && j < w) {
      const int x = x0 + i, y = y0 + j, n = x * ny + y;
      const int k = edge_code<K>(__ldg(graph + n), n, y, nx, ny);
      if (k >= 0) {
        const int ri = i + shift_dx(k), rj = j + shift_dy(k);
        if (ri >= 0 && ri < h && rj >= 0 && rj < w) {
          cd = (unsigned char)k;
          const int r = ri * T + rj;
          atomicOr(&donors32[r >> 2], (1u << k) << (8 * (r & 3)));
        } else {
          cd = (unsigned char)(k | 8);
        }
      }
      v = __ldg(src + n);
    }
    sm.code[c] = cd;
    sm.val[c] = v;
  }
  __syncthreads();

  if (FINAL) {
    // perimeter cells also receive the finished totals of other tiles' exit cells (a compact loop over the
    // perimeter: one lane per cell, instead of a divergent branch inside the load loop)
    for (int p = tid; p < 4 * T; p += kThreads) {
      int i, j;
      if (!perimeter_cell(p, h, w, i, j)) continue;
      const int x = x0 + i, y = y0 + j;
      sm.val[i * T + j] += external_inflow<K>(graph, out, x, y, x * ny + y, x0, y0, h, w, nx, ny);
    }
  }

  // ---- pending counts; leaves seed the first ready queue
#pragma unroll 4
  for (int m = 0; m < kCellsPerThread; ++m) {
    const int c = tid + m * kThreads, i = c / T, j = c % T;
    const unsigned mask = sm.donors[c];
    if (i < h && j < w && mask == 0) sm.queue[0][push_slot(&sm.count[0])] = (unsigned short)c;
    pend8[c] = (unsigned char)__popc(mask);
  }
  __syncthreads();

  // ---- rounds over the ready queue.  A walker finishes its cell (source + finished donors in slot order),
  // delivers it to the receiver, and carries on through the receiver only if it delivered the last pending
  // donor - for at most kWalkSteps cells, then the chain is re-queued so that no lane is stuck with a river.
  int cur = 0;
  while (true) {
    const int nready = sm.count[cur];
    if (nready == 0) break;
    // many ready cells: one cell per walker and re-queue, so that the lanes stay dense (most walkers stop at
    // their first junction anyway); few ready cells: follow the chains, fewer block-wide rounds
#ifndef SOILB200_WALK_DENSE
#define SOILB200_WALK_DENSE 1
#endif
#ifndef SOILB200_WALK_MID
#define SOILB200_WALK_MID 4
#endif
    const int wsteps = nready >= kThreads ? SOILB200_WALK_DENSE : (nready >= kThreads / 4 ? SOILB200_WALK_MID : kWalkSteps);
    for (int idx = tid; idx < nready; idx += kThreads) {
      int c = sm.queue[cur][idx];
      for (int step = 0;; ++step) {
        unsigned mask = sm.donors[c];
        if (mask) {
          float v = val[c];
          while (mask) {
            const int k = __ffs(mask) - 1;
            mask &= mask - 1;
            v += val[c - tile_offset(k)];
          }
          val[c] = v;
        }
        const unsigned char cd = sm.code[c];
        if (cd >= 8) break;  // leaves the tile, or no receiver
        const int r = c + tile_offset(cd);
        // a receiver with a single donor (most of a river's cells) has nobody else to wait for: no counter
        // traffic and no fences on that edge
        const unsigned rmask = sm.donors[r];
        if ((rmask & (rmask - 1u)) != 0u) {
          __threadfence_block();
          const unsigned sh = 8u * (r & 3);
          const unsigned before = atomicSub(&sm.pend32[r >> 2], 1u << sh);
          if (((before >> sh) & 0xFFu) != 1u) break;  // other donors pending: their last walker takes over
          __threadfence_block();
        }
        c = r;
        if (step + 1 >= wsteps) {
          sm.queue[cur ^ 1][push_slot(&sm.count[cur ^ 1])] = (unsigned short)c;
          break;
        }
      }
    }
    __syncthreads();
    if (tid == 0) sm.count[cur] = 0;
    cur ^= 1;
    __syncthreads();
  }

  if (!FINAL) {
    // ---- perimeter: which exit cell does each perimeter cell drain to; exit cells publish local totals
    for (int p = tid; p < 4 * T; p += kThreads) {
      int i, j;
      if (!perimeter_cell(p, h, w, i, j)) continue;
      int c = i * T + j;
      int steps = 0;
      while (sm.code[c] < 8 && steps++ < TT) c += tile_offset(sm.code[c]);
      const int n = (x0 + i) * ny + (y0 + j);
      const bool drains_out = sm.code[c] != kNone && sm.code[c] >= 8;
      root[n] = drains_out ? (x0 + c / T) * ny + (y0 + c % T) : -1;
      const unsigned char own = sm.code[i * T + j];
      if (own != kNone && own >= 8) {  // this perimeter cell is itself an exit cell
        out[n] = sm.val[i * T + j];
        pend2[n] = 0;
      }
    }
  } else {
    // ---- every cell except the exit cells (their totals were finished in phase B)
#pragma unroll 4
    for (int m = 0; m < kCellsPerThread; ++m) {
      const int c = tid + m * kThreads, i = c / T, j = c % T;
      if (i < h && j < w) {
        const unsigned char cd = sm.code[c];
        if (cd == kNone || cd < 8) out[(size_t)(x0 + i) * ny + (y0 + j)] = sm.val[c];
      }
    }
  }
}

// reduced forest: for each exit cell e, next[e] = exit cell that e's receiver drains to (or -1); count donors
template <int K>
__global__ void __launch_bounds__(kThreads) k_reduce_link(const int32_t* __restrict__ graph,
                                                          const int* __restrict__ root, int* pend2,
                                                          int* __restrict__ next, int nx, int ny, int tiles_y) {
  const int x0 = (blockIdx.x / tiles_y) * T, y0 = (blockIdx.x % tiles_y) * T;
  const int h = min(T, nx - x0), w = min(T, ny - y0);
  for (int p = threadIdx.x; p < 4 * T; p += kThreads) {
    int i, j;
    if (!perimeter_cell(p, h, w, i, j)) continue;
    const int x = x0 + i, y = y0 + j, n = x * ny + y;
    const int k = edge_code<K>(__ldg(graph + n), n, y, nx, ny);
    if (k < 0) continue;
    const int ri = i + shift_dx(k), rj = j + shift_dy(k);
    if (ri >= 0 && ri < h && rj >= 0 && rj < w) continue;  // not an exit cell
    const int r = (x + shift_dx(k)) * ny + (y + shift_dy(k));
    const int t = __ldg(root + r);
    next[n] = t;
    if (t >= 0) atomicAdd(pend2 + t, 1);
  }
}

// mode 0: mark reduced leaves (pending 0 -> -1, a value a decrement can never produce); mode 1: walk
template <int K, int MODE>
__global__ void __launch_bounds__(kThreads) k_reduce_walk(const int32_t* __restrict__ graph, int* pend2,
                                                          const int* __restrict__ next, float* out, int nx, int ny,
                                                          int tiles_y) {
  const int x0 = (blockIdx.x / tiles_y) * T, y0 = (blockIdx.x % tiles_y) * T;
  const int h = min(T, nx - x0), w = min(T, ny - y0);
  for (int p = threadIdx.x; p < 4 * T; p += kThreads) {
    int i, j;
    if (!perimeter_cell(p, h, w, i, j)) continue;
    const int x = x0 + i, y = y0 + j, n = x * ny + y;
    const int k = edge_code<K>(__ldg(graph + n), n, y, nx, ny);
    if (k < 0) continue;
    const int ri = i + shift_dx(k), rj = j + shift_dy(k);
    if (ri >= 0 && ri < h && rj >= 0 && rj < w) continue;  // not an exit cell
    if (MODE == 0) {
      if (pend2[n] == 0) pend2[n] = -1;
      continue;
    }
    if (__ldcg(pend2 + n) != -1) continue;
    int e = n;
    while (true) {
      const float v = __ldcg(out + e);
      const int t = __ldg(next + e);
      if (t < 0) break;
      atomicAdd(out + t, v);
      __threadfence();
      const int before = atomicSub(pend2 + t, 1);
      if (before != 1) break;
      __threadfence();
      e = t;
    }
  }
}

}  // namespace

size_t accumulate_tiled_ws(int nx, int ny) { return (size_t)nx * ny * 3 * sizeof(int); }

int launch_accumulate_tiled(const int32_t* graph, const float* src, float* out, int nx, int ny, int edge, void* ws,
                            cudaStream_t s) {
  const size_t N = (size_t)nx * ny;
  int* root = (int*)ws;
  int* pend2 = root + N;
  int* next = pend2 + N;
  const int tiles_x = (nx + T - 1) / T, tiles_y = (ny + T - 1) / T;
  const unsigned nb = (unsigned)(tiles_x * tiles_y);
  if (edge == SOILB200_D4) {
    k_tile<4, false><<<nb, kThreads, 0, s>>>(graph, src, out, root, pend2, nx, ny, tiles_y);
    k_reduce_link<4><<<nb, kThreads, 0, s>>>(graph, root, pend2, next, nx, ny, tiles_y);
    k_reduce_walk<4, 0><<<nb, kThreads, 0, s>>>(graph, pend2, next, out, nx, ny, tiles_y);
    k_reduce_walk<4, 1><<<nb, kThreads, 0, s>>>(graph, pend2, next, out, nx, ny, tiles_y);
    k_tile<4, true><<<nb, kThreads, 0, s>>>(graph, src, out, root, pend2, nx, ny, tiles_y);
  } else {
    k_tile<8, false><<<nb, kThreads, 0, s>>>(graph, src, out, root, pend2, nx, ny, tiles_y);
    k_reduce_link<8><<<nb, kThreads, 0, s>>>(graph, root, pend2, next, nx, ny, tiles_y);
    k_reduce_walk<8, 0><<<nb, kThreads, 0, s>>>(graph, pend2, next, out, nx, ny, tiles_y);
    k_reduce_walk<8, 1><<<nb, kThreads, 0, s>>>(graph, pend2, next, out, nx, ny, tiles_y);
    k_tile<8, true><<<nb, kThreads, 0, s>>>(graph, src, out, root, pend2, nx, ny, tiles_y);
  }
  count_launch(5);
  return check_launch();
}

}  // namespace sb
