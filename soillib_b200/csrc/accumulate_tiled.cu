// accumulate_tiled.cu - upstream accumulation (with or without per-edge decay) resolved tile by tile in shared memory.
//
// Reference: source/soillib/model/graph/graph.cu:321-593 (soil::accumulate, soil::accumulate_decay).  Same sums as
// accumulate.cu's global walk, but the long serial chains (a river is tens of thousands of cells) are resolved inside
// 64x64 tiles at shared-memory latency, and only tile-boundary cells meet in global memory.  Accumulation is LINEAR in
// the sources (value[n] = source[n] + sum_k w_k * value[donor_k], graph.cu:436-520), which is what the phases use:
//
//   A  per tile (k_exits): what leaves the tile.  Every cell finds the last cell of its in-tile path - an "exit" cell
//      (receiver in another tile) or a sink - by pointer jumping on the receiver pointers (log2(path length) rounds, all
//      cells in parallel, no ordering), carrying the product of the edge weights along; then source * product is added to
//      that cell's total.  Exit cells publish their in-tile totals; every perimeter cell records the exit it drains to
//      and the weight product on the way.  No cell is walked in order here: this phase used to be the same ordered walk
//      as phase C (1.3 of accumulate's 2.7 ms at 8192^2).
//   B  reduced forest over exit cells only: exit e -> the exit that e's receiver drains to, with the weight of the edge
//      e -> receiver times the receiver's product.  Counter walk as in accumulate.cu, totals added with float atomics.
//   C  per tile (k_tile): perimeter cells start from source + the finished totals of the exit cells that flow into them
//      from outside, walkers resolve the tile in order (donors in the reference's slot order), every non-exit cell is
//      written.
//
// Per-edge decay (my_decay, graph.cu:382-420): the edge d -> r carries decay[d] if d sits in one of the first four
// slots of r's compacted donor list, else __powf(decay[d], 1.414f); the slot of d is the number of r's donors with a
// smaller neighbour code.  For a perimeter cell that count includes donors in other tiles.
//
// HBM traffic: A reads graph + source (+ decay), C reads them again and writes out; B touches perimeter cells only.
// Scratch 12 B/cell (root, pending, next), 20 B/cell with decay (+ product to the root, + weight of the reduced edge).
// Sums of integer-valued sources are exact; real-valued sources differ from the global walk in summation order only.
#include "common.cuh"

namespace sb {
namespace {

constexpr int T = 64;               // tile edge
constexpr int TT = T * T;
#ifndef SOILB200_TILE_THREADS
#define SOILB200_TILE_THREADS 512
#endif
constexpr int kThreads = SOILB200_TILE_THREADS;
constexpr int kCellsPerThread = TT / kThreads;  // 8
constexpr unsigned char kNone = 0xFF;

// graph.hpp:35-44 without the constant-bank tables (an indexed constant load serialises over the distinct k of a
// warp): the x and y components of shift_k, two bits each, biased by one
__device__ __forceinline__ int shift_dx(int k) { return (int)((0xA094u >> (2 * k)) & 3u) - 1; }
__device__ __forceinline__ int shift_dy(int k) { return (int)((0x8861u >> (2 * k)) & 3u) - 1; }

// neighbour code k of the edge n -> r, or -1 if r is not an in-map K-neighbour of n (graph.cu:336-346); dx, dy = shift_k.
// Without branches (the chain of comparisons it replaces split every warp eight ways: a fifth of both tile kernels'
// instructions): a = r - n + ny + 1 is the position of r in the 3x3 block around n, rows ny apart - unambiguous for
// ny >= 3 (the launcher sends narrower maps to the global walk).
template <int K>
__device__ __forceinline__ int edge_code(int r, int n, int y, int nx, int ny, int& dx, int& dy) {
  const int a = r - n + ny + 1;
  const int bx = (a >= ny) + (a >= 2 * ny);   // 0, 1, 2 <-> dx = -1, 0, 1
  const int by = a - bx * ny;                 // 0, 1, 2 <-> dy = -1, 0, 1
  bool ok = (unsigned)r < (unsigned)(nx * ny) && (unsigned)by <= 2u;
  ok = ok && !(by == 0 && y == 0) && !(by == 2 && y == ny - 1);  // would wrap around a row end
  // (bx, by) -> k of graph.hpp:35-44, four bits each; the centre is no neighbour
  const int k = (int)((0x7362F1504ull >> (4 * (ok ? bx * 3 + by : 4))) & 0xFull);
  dx = bx - 1; dy = by - 1;
  return k < K ? k : -1;
}
template <int K>
__device__ __forceinline__ int edge_code(int r, int n, int y, int nx, int ny) {
  int dx, dy;
  return edge_code<K>(r, n, y, nx, ny, dx, dy);
}

// shift_k as an offset inside the T x T tile (graph.hpp:35-44: -T, -1, 1, T, -T-1, -T+1, T-1, T+1), one signed byte
// per k packed in a 64-bit constant: two ALU instructions instead of an indexed constant-bank load
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

// my_decay, graph.cu:396-413: the weight of the edge d -> r given the slot d occupies in r's compacted donor list
__device__ __forceinline__ float edge_weight(float decay_d, int slot) { return slot < 4 ? decay_d : __powf(decay_d, 1.414f); }
__device__ __forceinline__ int slot_of(unsigned full_mask, int k) { return __popc(full_mask & ((1u << k) - 1u)); }

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

// bit k set: the cell at (x,y) - shift[k] lies in ANOTHER tile (inside the map) and flows into (x,y)
template <int K>
__device__ __forceinline__ unsigned foreign_donors(const int32_t* __restrict__ graph, int x, int y, int x0, int y0, int h,
                                                   int w, int nx, int ny) {
  unsigned m = 0u;
#pragma unroll
  for (int k = 0; k < K; ++k) {  // slot order: the donor that reaches n by shift k sits at n - shift[k]
    const int qx = x - c_shift_x[k], qy = y - c_shift_y[k];
    if (qx < 0 || qy < 0 || qx >= nx || qy >= ny) continue;
    if (qx >= x0 && qx < x0 + h && qy >= y0 && qy < y0 + w) continue;  // in-tile donors are in the tile's masks
    const int q = qx * ny + qy;
    if (edge_code<K>(__ldg(graph + q), q, qy, nx, ny) == k) m |= 1u << k;
  }
  return m;
}
// all donors of the cell (x,y), in any tile (the reference's __donor, graph.cu:336-346, seen from the receiver)
template <int K>
__device__ __forceinline__ unsigned all_donors(const int32_t* __restrict__ graph, int x, int y, int nx, int ny) {
  unsigned m = 0u;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int qx = x - c_shift_x[k], qy = y - c_shift_y[k];
    if (qx < 0 || qy < 0 || qx >= nx || qy >= ny) continue;
    const int q = qx * ny + qy;
    if (edge_code<K>(__ldg(graph + q), q, qy, nx, ny) == k) m |= 1u << k;
  }
  return m;
}

// what both tile kernels keep per cell
template <bool FULL>
struct tile_graph {
  unsigned char code[TT];                 // 0..7 receiver inside the tile, 8..15 receiver in another tile, 0xFF none
  unsigned char donors[FULL ? TT : 4];    // bit k set: the in-tile cell at c - shift[k] flows into c
  unsigned char fdon[FULL ? TT : 4];      // (decay) bit k set: the cell at c - shift[k] of ANOTHER tile flows into c (perimeter only)
};

// receiver codes and donor masks of the tile; with WEIGHTS also the weight of every cell's in-tile edge (wgt[c], 1 if
// the cell has no in-tile receiver).  Ends with a barrier.
template <int K, bool WEIGHTS, bool DONORS = true>
__device__ __forceinline__ void load_graph(tile_graph<DONORS>& g, float* wgt, const int32_t* __restrict__ graph,
                                           const float* __restrict__ decay, int x0, int y0, int h, int w, int nx, int ny) {
  static_assert(DONORS || !WEIGHTS, "edge weights need the donor masks");
  const int tid = threadIdx.x;
  unsigned* donors32 = reinterpret_cast<unsigned*>(g.donors);
  unsigned* fdon32 = reinterpret_cast<unsigned*>(g.fdon);
  if (DONORS) {
    for (int c4 = tid; c4 < TT / 4; c4 += kThreads) { donors32[c4] = 0u; if (WEIGHTS) fdon32[c4] = 0u; }
    __syncthreads();
  }
#pragma unroll 4
  for (int m = 0; m < kCellsPerThread; ++m) {
    const int c = tid + m * kThreads, i = c / T, j = c % T;
    unsigned char cd = kNone;
    if (i < h && j < w) {
      const int x = x0 + i, y = y0 + j, n = x * ny + y;
      int dx, dy;
      const int k = edge_code<K>(__ldg(graph + n), n, y, nx, ny, dx, dy);
      if (k >= 0) {
        const int ri = i + dx, rj = j + dy;
        if (ri >= 0 && ri < h && rj >= 0 && rj < w) {
          cd = (unsigned char)k;
          const int r = ri * T + rj;
          if (DONORS) atomicOr(&donors32[r >> 2], (1u << k) << (8 * (r & 3)));
        } else {
          cd = (unsigned char)(k | 8);
        }
      }
    }
    g.code[c] = cd;
  }
  if (WEIGHTS) {
    // the slot of a donor counts the receiver's donors in other tiles too: perimeter cells look them up
    for (int p = tid; p < 4 * T; p += kThreads) {
      int i, j;
      if (!perimeter_cell(p, h, w, i, j)) continue;
      g.fdon[i * T + j] = (unsigned char)foreign_donors<K>(graph, x0 + i, y0 + j, x0, y0, h, w, nx, ny);
    }
  }
  __syncthreads();
  if (WEIGHTS) {
#pragma unroll 4
    for (int m = 0; m < kCellsPerThread; ++m) {
      const int c = tid + m * kThreads, i = c / T, j = c % T;
      float wv = 1.0f;
      const unsigned char cd = g.code[c];
      if (i < h && j < w && cd < 8) {
        const int r = c + tile_offset(cd);
        const unsigned full = (unsigned)g.donors[r] | (unsigned)g.fdon[r];
        wv = edge_weight(__ldg(decay + (size_t)(x0 + i) * ny + (y0 + j)), slot_of(full, cd));
      }
      wgt[c] = wv;
    }
    __syncthreads();
  }
}

// ---- phase A: exit totals by pointer jumping --------------------------------------------------------------------
template <bool DECAY>
struct exits_smem {
  tile_graph<DECAY> g;
  unsigned short ptr[TT];          // a cell further down the in-tile path; converges to its last cell
  float tot[TT];                   // sum over the cells that end here of source * weight product
  float mul[DECAY ? TT : 4];       // (decay) product of the edge weights from the cell to ptr[cell]
};

template <int K, bool DECAY>
__global__ void __launch_bounds__(kThreads, 2048 / kThreads) k_exits(const int32_t* __restrict__ graph, const float* __restrict__ src,
                                                    const float* __restrict__ decay, float* out, int* __restrict__ root,
                                                    float* __restrict__ mroot, int* __restrict__ pend2, int nx, int ny,
                                                    int tiles_y) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  exits_smem<DECAY>& sm = *reinterpret_cast<exits_smem<DECAY>*>(smem_raw);
  const int x0 = (blockIdx.x / tiles_y) * T, y0 = (blockIdx.x % tiles_y) * T;
  const int h = min(T, nx - x0), w = min(T, ny - y0);
  const int tid = threadIdx.x;

  load_graph<K, DECAY, DECAY>(sm.g, sm.mul, graph, decay, x0, y0, h, w, nx, ny);
#pragma unroll 4
  for (int m = 0; m < kCellsPerThread; ++m) {
    const int c = tid + m * kThreads;
    const unsigned char cd = sm.g.code[c];
    sm.ptr[c] = (unsigned short)(cd < 8 ? c + tile_offset(cd) : c);
    sm.tot[c] = 0.0f;
  }
  __syncthreads();

  // ptr <- ptr[ptr], mul <- mul * mul[ptr]: after round i a cell points 2^i cells down its path (or at its end).
  // A cell whose pointer has reached the end of its path drops out (most paths are short; the rounds go on for the
  // longest one).  13 rounds cover any path of a 4096-cell tile; a cyclic (invalid) graph just stops there.
  unsigned active = (1u << kCellsPerThread) - 1u;
  for (int round = 0; round < 13; ++round) {
    unsigned short pp[kCellsPerThread];
    float mp[kCellsPerThread];
    unsigned moved = 0u;
#pragma unroll
    for (int m = 0; m < kCellsPerThread; ++m) {
      if (!((active >> m) & 1u)) continue;
      const int c = tid + m * kThreads;
      const int p = sm.ptr[c];
      pp[m] = sm.ptr[p];
      if (DECAY) mp[m] = sm.mul[p];
      if (pp[m] != p) moved |= 1u << m;   // p was not yet an end: take its pointer and its product
    }
    active = moved;
    __syncthreads();
#pragma unroll
    for (int m = 0; m < kCellsPerThread; ++m) {
      if (!((moved >> m) & 1u)) continue;
      const int c = tid + m * kThreads;
      sm.ptr[c] = pp[m];
      if (DECAY) sm.mul[c] = sm.mul[c] * mp[m];
    }
    if (!__syncthreads_or((int)moved)) break;
  }

  // a warp holds 32 neighbouring cells of one row, which mostly end in the same few cells: runs of lanes with the same
  // end add up among themselves first (segmented scan, five shuffle steps, always in lane order) and the last lane of a
  // run posts ONE shared-memory atomic - hundreds of lanes adding to one address one by one (a float add in shared
  // memory is a compare-and-swap loop) cost 2 ms
  const int lane = tid & 31;
  for (int m = 0; m < kCellsPerThread; ++m) {
    const int c = tid + m * kThreads, i = c / T, j = c % T;
    const bool in = i < h && j < w;
    float v = 0.0f;
    if (in) {
      v = __ldg(src + (size_t)(x0 + i) * ny + (y0 + j));
      if (DECAY) v = v * sm.mul[c];
    }
    const int end = in ? (int)sm.ptr[c] : -1 - lane;   // cells outside the map never share a run
    const int prev = __shfl_up_sync(0xffffffffu, end, 1);
    const unsigned heads = __ballot_sync(0xffffffffu, lane == 0 || prev != end);
    const int start = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));   // first lane of this lane's run
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const float o = __shfl_up_sync(0xffffffffu, v, d);
      if (lane - d >= start) v += o;
    }
    const bool last = lane == 31 || ((heads >> (lane + 1)) & 1u);
    if (in && last) atomicAdd(&sm.tot[end], v);
  }
  __syncthreads();

  // ---- perimeter: which exit cell does each perimeter cell drain to; exit cells publish their in-tile totals
  for (int p = tid; p < 4 * T; p += kThreads) {
    int i, j;
    if (!perimeter_cell(p, h, w, i, j)) continue;
    const int c = i * T + j, e = sm.ptr[c];
    const int n = (x0 + i) * ny + (y0 + j);
    const unsigned char ce = sm.g.code[e];
    const bool drains_out = ce != kNone && ce >= 8;
    root[n] = drains_out ? (x0 + e / T) * ny + (y0 + e % T) : -1;
    if (DECAY) mroot[n] = sm.mul[c];
    const unsigned char own = sm.g.code[c];
    if (own != kNone && own >= 8) {  // this perimeter cell is itself an exit cell
      out[n] = sm.tot[c];
      pend2[n] = 0;
    }
  }
}

// ---- phase C: the ordered walk --------------------------------------------------------------------------------
// One 8-byte record per cell - the value and a word with everything a walker needs to know about the cell - so that a
// step along a river is ONE shared-memory round trip and a dozen instructions (the walk is what a tile waits for: a
// steepest-descent graph on smooth terrain is a few hundred parallel flow lines, each as long as the tile).
//   word: bits 0-11 in-tile receiver, bit 12 has one, bit 13 receiver in another tile (an exit cell),
//         bits 16-23 in-tile donor mask (bit k: the cell at c - shift[k] flows into c), bits 24-27 donors pending,
//         bits 28-31 neighbour code of the in-tile receiver
constexpr unsigned kIn = 1u << 12, kExit = 1u << 13;
// threads per tile and tiles per SM: without decay a tile takes 40 KB and five of them fit an SM - 256 threads each, the
// walk leaves most lanes idle anyway and a fifth tile fills the issue slots; with decay (60 KB) three tiles of 512
__host__ __device__ constexpr int walk_threads(bool decay) { return decay ? 512 : 256; }
__host__ __device__ constexpr int walk_blocks(bool decay) { return decay ? 3 : 5; }

template <bool DECAY>
struct tile_smem {
  uint2 cell[TT];               // .x value (float bits), .y word
  unsigned short queue[TT];     // ready cells, round after round (a cell enters at most once)
  int count[2];
  float wgt[1];                 // (decay only; never touched here)
  unsigned char fdon[1];
};
template <>
struct tile_smem<true> {
  uint2 cell[TT];
  unsigned short queue[TT];
  int count[2];
  float wgt[TT];                // weight of the cell's in-tile edge
  unsigned char fdon[TT];       // bit k set: the cell at c - shift[k] of ANOTHER tile flows into c (perimeter only)
};

template <int K, bool DECAY>
__global__ void __launch_bounds__(walk_threads(DECAY), walk_blocks(DECAY)) k_tile(const int32_t* __restrict__ graph, const float* __restrict__ src,
                                                   const float* __restrict__ decay, float* out, int nx, int ny,
                                                   int tiles_y) {
  constexpr int kWalkThreads = walk_threads(DECAY), kWalkCells = TT / kWalkThreads;
  static_assert(kWalkThreads >= 4 * T, "one lane per perimeter cell");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  tile_smem<DECAY>& sm = *reinterpret_cast<tile_smem<DECAY>*>(smem_raw);
  const int x0 = (blockIdx.x / tiles_y) * T, y0 = (blockIdx.x % tiles_y) * T;
  const int h = min(T, nx - x0), w = min(T, ny - y0);
  const int tid = threadIdx.x;

  if (tid < 2) sm.count[tid] = 0;
#pragma unroll
  for (int m = 0; m < kWalkCells; ++m) {
    sm.cell[tid + m * kWalkThreads].y = 0u;
    if (DECAY) sm.fdon[tid + m * kWalkThreads] = 0;
  }

  // ---- one lane per perimeter cell: the finished totals of the other tiles' exit cells that flow into it, summed in
  // the reference's slot order (the slots count the cell's in-tile donors too).  First, so that these dependent loads
  // (neighbour's receiver, then its total) are in flight under the tile's own loads.
  int pc = -1;
  float inflow = 0.0f;
  {
    int i, j;
    if (tid < 4 * T && perimeter_cell(tid, h, w, i, j)) {
      const int x = x0 + i, y = y0 + j;
      unsigned full = 0u, fm = 0u;
#pragma unroll
      for (int k = 0; k < K; ++k) {  // slot order: the donor that reaches n by shift k sits at n - shift[k]
        const int qx = x - c_shift_x[k], qy = y - c_shift_y[k];
        if (qx < 0 || qy < 0 || qx >= nx || qy >= ny) continue;
        const bool inside = qx >= x0 && qx < x0 + h && qy >= y0 && qy < y0 + w;
        if (!DECAY && inside) continue;  // without decay the slots do not matter
        const int q = qx * ny + qy;
        if (edge_code<K>(__ldg(graph + q), q, qy, nx, ny) != k) continue;
        full |= 1u << k;
        if (!inside) fm |= 1u << k;
      }
      if (fm) {
        pc = i * T + j;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          if (!((fm >> k) & 1u)) continue;
          const int q = (x - c_shift_x[k]) * ny + (y - c_shift_y[k]);
          if (DECAY) inflow += edge_weight(__ldg(decay + q), slot_of(full, k)) * __ldcg(out + q);
          else inflow += __ldcg(out + q);
        }
        if (DECAY) sm.fdon[pc] = (unsigned char)fm;
      }
    }
  }
  __syncthreads();  // words zeroed

  // ---- the tile's cells: value, receiver, donor masks (four of a lane's cells at a time: registers)
#pragma unroll
  for (int half = 0; half < kWalkCells / 4; ++half) {
    constexpr int H = 4;
    int gr[H];
    float sv[H], dq[H];
#pragma unroll
    for (int m = 0; m < H; ++m) {
      const int c = tid + (half * H + m) * kWalkThreads, i = c / T, j = c % T;
      const bool in = i < h && j < w;
      const size_t n = (size_t)(x0 + i) * ny + (y0 + j);
      gr[m] = in ? __ldg(graph + n) : -1;
      sv[m] = in ? __ldg(src + n) : 0.0f;
      if (DECAY) dq[m] = in ? __ldg(decay + n) : 1.0f;
    }
#pragma unroll
    for (int m = 0; m < H; ++m) {
      const int c = tid + (half * H + m) * kWalkThreads, i = c / T, j = c % T;
      unsigned base = 0u;
      if (i < h && j < w) {
        int dx, dy;
        const int n = (x0 + i) * ny + (y0 + j);
        const int k = edge_code<K>(gr[m], n, y0 + j, nx, ny, dx, dy);
        if (k >= 0) {
          const int ri = i + dx, rj = j + dy;
          if (ri >= 0 && ri < h && rj >= 0 && rj < w) {
            const int r = ri * T + rj;
            base = (unsigned)r | kIn | ((unsigned)k << 28);
            atomicOr(&sm.cell[r].y, 1u << (16 + k));
          } else {
            base = kExit;
          }
        }
      }
      if (base) atomicOr(&sm.cell[c].y, base);   // donors are adding their bits to the same word
      sm.cell[c].x = __float_as_uint(sv[m]);
      if (DECAY) sm.wgt[c] = dq[m];              // the cell's decay for now; its edge weight once the masks are complete
    }
  }
  __syncthreads();
  if (pc >= 0) sm.cell[pc].x = __float_as_uint(__uint_as_float(sm.cell[pc].x) + inflow);

  // ---- pending counts, edge weights; leaves seed the first ready queue (one reservation per warp: a lane's leaves
  // are counted, the counts scanned over the warp, and lane 31 takes the warp's share of the queue)
  unsigned leaves = 0u;
#pragma unroll
  for (int m = 0; m < kWalkCells; ++m) {
    const int c = tid + m * kWalkThreads, i = c / T, j = c % T;
    const unsigned wd = sm.cell[c].y;
    const unsigned mask = (wd >> 16) & 0xFFu;
    if (mask) sm.cell[c].y = wd | ((unsigned)__popc(mask) << 24);
    if (i < h && j < w && mask == 0u) leaves |= 1u << m;
    if (DECAY) {
      float wv = 1.0f;
      if (wd & kIn) {
        const int r = (int)(wd & 0xFFFu);
        const unsigned full = ((sm.cell[r].y >> 16) & 0xFFu) | (unsigned)sm.fdon[r];
        wv = edge_weight(sm.wgt[c], slot_of(full, (int)(wd >> 28)));
      }
      sm.wgt[c] = wv;
    }
  }
  {
    const int lane = tid & 31, mine = __popc(leaves);
    int upto = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, upto, d);
      if (lane >= d) upto += o;
    }
    int base = 0;
    if (lane == 31) base = atomicAdd(&sm.count[0], upto);
    int slot = __shfl_sync(0xffffffffu, base, 31) + upto - mine;
#pragma unroll
    for (int m = 0; m < kWalkCells; ++m)
      if ((leaves >> m) & 1u) sm.queue[slot++] = (unsigned short)(tid + m * kWalkThreads);
  }
  __syncthreads();

  // ---- rounds over the ready queue.  A walker finishes its cell (source + finished donors in slot order),
  // delivers it to the receiver, and carries on through the receiver only if it delivered the last pending
  // donor.  While the frontier is wider than the block a walker finishes one cell per round and re-queues its
  // chain (dense lanes); after that every ready chain has a lane of its own and is walked to its end, the
  // running value in a register along a river (one donor per cell: no counter, no fence, no reload).
  // volatile: other walkers' stores must be seen after the counter handshake; a record is one 8-byte load
  const volatile unsigned long long* rec = reinterpret_cast<const volatile unsigned long long*>(sm.cell);
  volatile float* val = reinterpret_cast<volatile float*>(sm.cell);   // val[2 * c]: the value of cell c
  int cur = 0, head = 0;
  while (true) {
    const int nready = sm.count[cur];
    if (nready == 0) break;
    const bool dense = nready >= kWalkThreads;
    for (int idx = tid; idx < nready; idx += kWalkThreads) {
      int c = sm.queue[head + idx];
      const unsigned long long first = rec[c];
      float v = __uint_as_float((unsigned)first);
      unsigned wd = (unsigned)(first >> 32);
      unsigned mask = (wd >> 16) & 0xFFu;
      while (true) {
        if (mask) {   // the finished donors, in the reference's slot order
#pragma unroll
          for (int k = 0; k < K; ++k) {
            if (!((mask >> k) & 1u)) continue;
            const int d = c - tile_offset(k);
            if (DECAY) v += sm.wgt[d] * val[2 * d];
            else v += val[2 * d];
          }
          val[2 * c] = v;
        }
        if (!(wd & kIn)) break;  // leaves the tile, or no receiver
        const int r = (int)(wd & 0xFFFu);
        const unsigned long long rc = rec[r];   // the receiver's own term (only the walker that finishes r writes it) and word
        const unsigned rwd = (unsigned)(rc >> 32);
        const float rv = __uint_as_float((unsigned)rc);
        const unsigned rmask = (rwd >> 16) & 0xFFu;
        const bool shared = (rmask & (rmask - 1u)) != 0u;
        if (shared) {
          __threadfence_block();
          const unsigned before = atomicSub(&sm.cell[r].y, 1u << 24);
          if (((before >> 24) & 0xFu) != 1u) break;  // other donors pending: their last walker takes over
          __threadfence_block();
        }
        if (dense) {
          sm.queue[head + nready + push_slot(&sm.count[cur ^ 1])] = (unsigned short)r;
          break;
        }
        if (shared) {
          mask = rmask; v = rv;
        } else {
          mask = 0u;
          v = DECAY ? rv + sm.wgt[c] * v : rv + v;
          val[2 * r] = v;
        }
        c = r; wd = rwd;
      }
    }
    __syncthreads();
    if (tid == 0) sm.count[cur] = 0;
    head += nready;
    cur ^= 1;
    __syncthreads();
  }

  // ---- every cell except the exit cells (their totals were finished in phase B)
#pragma unroll
  for (int m = 0; m < kWalkCells; ++m) {
    const int c = tid + m * kWalkThreads, i = c / T, j = c % T;
    if (i < h && j < w) {
      const uint2 cc = sm.cell[c];
      if (!(cc.y & kExit)) out[(size_t)(x0 + i) * ny + (y0 + j)] = __uint_as_float(cc.x);
    }
  }
}

// ---- phase B: the reduced forest ------------------------------------------------------------------------------
// for each exit cell e: next[e] = exit cell that e's receiver drains to (or -1), wnext[e] = weight of the edge
// e -> receiver times the product from the receiver to that exit; count donors
template <int K, bool DECAY>
__global__ void __launch_bounds__(kThreads) k_reduce_link(const int32_t* __restrict__ graph,
                                                          const float* __restrict__ decay, const int* __restrict__ root,
                                                          const float* __restrict__ mroot, int* pend2,
                                                          int* __restrict__ next, float* __restrict__ wnext, int nx,
                                                          int ny, int tiles_y) {
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
    const int rx = x + shift_dx(k), ry = y + shift_dy(k);
    const int r = rx * ny + ry;
    const int t = __ldg(root + r);
    next[n] = t;
    if (t >= 0) {
      if (DECAY) wnext[n] = edge_weight(__ldg(decay + n), slot_of(all_donors<K>(graph, rx, ry, nx, ny), k)) * __ldg(mroot + r);
      atomicAdd(pend2 + t, 0x10001);   // pending | ever (k_reduce_walk)
    }
  }
}

// one walker per reduced leaf (an exit cell no other exit drains to).  pend2 holds the donors still pending in its low
// half and, in its high half, the number of donors the cell ever had (k_reduce_link adds 0x10001 per donor): the high
// half tells a leaf from a cell whose donors have all arrived, so no marking pass is needed
template <int K, bool DECAY>
__global__ void __launch_bounds__(kThreads) k_reduce_walk(const int32_t* __restrict__ graph, int* pend2,
                                                          const int* __restrict__ next, const float* __restrict__ wnext,
                                                          float* out, int nx, int ny, int tiles_y) {
  const int x0 = (blockIdx.x / tiles_y) * T, y0 = (blockIdx.x % tiles_y) * T;
  const int h = min(T, nx - x0), w = min(T, ny - y0);
  for (int p = threadIdx.x; p < 4 * T; p += kThreads) {
    int i, j;
    if (!perimeter_cell(p, h, w, i, j)) continue;
    const int x = x0 + i, y = y0 + j, n = x * ny + y;
    int dx, dy;
    const int k = edge_code<K>(__ldg(graph + n), n, y, nx, ny, dx, dy);
    if (k < 0) continue;
    const int ri = i + dx, rj = j + dy;
    if (ri >= 0 && ri < h && rj >= 0 && rj < w) continue;  // not an exit cell
    if ((__ldcg(pend2 + n) >> 16) != 0) continue;          // has donors: their last walker comes through
    int e = n;
    while (true) {
      const float v = __ldcg(out + e);
      const int t = __ldg(next + e);
      if (t < 0) break;
      atomicAdd(out + t, DECAY ? __ldg(wnext + e) * v : v);
      __threadfence();
      const int before = atomicSub(pend2 + t, 1);
      if ((before & 0xFFFF) != 1) break;
      __threadfence();
      e = t;
    }
  }
}

template <int K, bool DECAY>
int launch_phases(const int32_t* graph, const float* src, const float* decay, float* out, int nx, int ny, void* ws,
                  cudaStream_t s) {
  const size_t N = (size_t)nx * ny;
  int* root = (int*)ws;
  int* pend2 = root + N;
  int* next = pend2 + N;
  float* mroot = DECAY ? (float*)(next + N) : nullptr;
  float* wnext = DECAY ? mroot + N : nullptr;
  const int tiles_x = (nx + T - 1) / T, tiles_y = (ny + T - 1) / T;
  const unsigned nb = (unsigned)(tiles_x * tiles_y);
  // both tile kernels keep more than the 48 KB a kernel gets without asking
  cudaError_t e = cudaFuncSetAttribute(k_exits<K, DECAY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(exits_smem<DECAY>));
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(k_tile<K, DECAY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(tile_smem<DECAY>));
  if (e != cudaSuccess) return (int)e;
  k_exits<K, DECAY><<<nb, kThreads, sizeof(exits_smem<DECAY>), s>>>(graph, src, decay, out, root, mroot, pend2, nx, ny, tiles_y);
  k_reduce_link<K, DECAY><<<nb, kThreads, 0, s>>>(graph, decay, root, mroot, pend2, next, wnext, nx, ny, tiles_y);
  k_reduce_walk<K, DECAY><<<nb, kThreads, 0, s>>>(graph, pend2, next, wnext, out, nx, ny, tiles_y);
  k_tile<K, DECAY><<<nb, walk_threads(DECAY), sizeof(tile_smem<DECAY>), s>>>(graph, src, decay, out, nx, ny, tiles_y);
  count_launch(4);
  return check_launch();
}

}  // namespace

size_t accumulate_tiled_ws(int nx, int ny, bool decay) { return (size_t)nx * ny * (decay ? 5 : 3) * sizeof(int); }

int launch_accumulate_tiled(const int32_t* graph, const float* src, const float* decay, float* out, int nx, int ny,
                            int edge, void* ws, cudaStream_t s) {
  if (edge == SOILB200_D4) {
    return decay ? launch_phases<4, true>(graph, src, decay, out, nx, ny, ws, s)
                 : launch_phases<4, false>(graph, src, nullptr, out, nx, ny, ws, s);
  }
  return decay ? launch_phases<8, true>(graph, src, decay, out, nx, ny, ws, s)
               : launch_phases<8, false>(graph, src, nullptr, out, nx, ny, ws, s);
}

}  // namespace sb
