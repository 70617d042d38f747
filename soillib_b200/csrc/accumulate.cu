// accumulate.cu - upstream accumulation of a source field along a receiver forest.
//
// Reference: source/soillib/model/graph/graph.cu:321-593 (donor lists, per-edge
// decay, 2*(ceil(log2(N)/2)+1) full-grid rake-compress rounds over
// 2*N*(8+8K) bytes of scratch).
//
// Same result, different algorithm.  value[n] = source[n] + sum_k w_k * value[donor_k],
// donors taken in neighbour order k = 0..K-1, which is exactly the compacted
// slot order of the reference's __count (graph.cu:350-380); w_k follows
// my_decay (graph.cu:382-420): the first four donors of a cell carry decay[d],
// later ones __powf(decay[d], 1.414f) (keyed on the slot).
// An edge n -> graph[n] exists only if graph[n] is one of the K in-bounds
// neighbours of n (__donor, graph.cu:336-346).
//
// Pass 1 counts donors per cell (radius-1 stencil on the int graph).  Pass 2
// starts one walker per leaf; a walker finishes a cell, publishes it, and
// decrements the receiver's pending counter - whoever brings it to zero owns
// the receiver and continues downstream, gathering the finished donors in
// fixed order.  Every cell is computed once, by one thread, in a fixed
// summation order: the result is deterministic and equals the exact
// topological sum.  Scratch: one int32 per cell.
#include "common.cuh"

namespace sb {

size_t accumulate_tiled_ws(int nx, int ny, bool decay);
int launch_accumulate_tiled(const int32_t* graph, const float* src, const float* decay, float* out, int nx, int ny,
                            int edge, void* ws, cudaStream_t s);

namespace {

// number of in-bounds neighbours whose receiver is n, and a bit mask of them
template <int K>
__device__ __forceinline__ unsigned donor_mask(const int32_t* __restrict__ graph, int x, int y, int nx, int ny) {
  const int n = x * ny + y;
  unsigned m = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int qx = x + c_shift_x[k], qy = y + c_shift_y[k];
    if (qx < 0 || qy < 0 || qx >= nx || qy >= ny) continue;
    if (__ldg(graph + (size_t)qx * ny + qy) == n) m |= 1u << k;
  }
  return m;
}

template <int K>
__global__ void __launch_bounds__(256) k_indegree(const int32_t* __restrict__ graph, const float* __restrict__ src,
                                                  int* __restrict__ pending, float* __restrict__ out, int nx, int ny) {
  const int y = blockIdx.x * 32 + (threadIdx.x & 31);
  const int x = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= nx || y >= ny) return;
  const size_t n = (size_t)x * ny + y;
  const int cnt = __popc(donor_mask<K>(graph, x, y, nx, ny));
  pending[n] = cnt == 0 ? -1 : cnt;  // -1 marks a leaf; a decrement can never produce it
  out[n] = src[n];  // cells on a cycle are never finished; they keep their own source
}

// receiver of n if it is a registered edge (graph[n] is an in-bounds K-neighbour), else -1
template <int K>
__device__ __forceinline__ int receiver_of(const int32_t* __restrict__ graph, int n, int nx, int ny) {
  const int r = __ldg(graph + n);
  if (r < 0 || r >= nx * ny) return -1;
  const int x = n / ny, y = n % ny, rx = r / ny, ry = r % ny;
  const int dx = rx - x, dy = ry - y;
  if (dx < -1 || dx > 1 || dy < -1 || dy > 1 || (dx == 0 && dy == 0)) return -1;
  if (K == 4 && dx != 0 && dy != 0) return -1;
  return r;
}

template <int K, bool DECAY>
__global__ void __launch_bounds__(256) k_walk(const int32_t* __restrict__ graph, const float* __restrict__ src,
                                              const float* __restrict__ decay, int* pending, float* out, int nx,
                                              int ny) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)nx * ny) return;
  int n = (int)t;
  // a walker is born on every leaf; pending was written by the previous launch
  if (__ldcg(pending + n) != -1) return;
  bool leaf = true;
  while (true) {
    const int x = n / ny, y = n % ny;
    float v = __ldg(src + n);
    if (!leaf) {
      int slot = 0;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        // slot k of the reference's donor list holds the cell that reaches n by shift k,
        // i.e. the neighbour at n - shift[k] (__donor, graph.cu:338-345)
        const int qx = x - c_shift_x[k], qy = y - c_shift_y[k];
        if (qx < 0 || qy < 0 || qx >= nx || qy >= ny) continue;
        const int d = qx * ny + qy;
        if (__ldg(graph + d) != n) continue;
        const float dv = __ldcg(out + d);  // finished by another walker: read at L2
        if (DECAY) {
          const float din = __ldg(decay + d);
          const float w = (slot < 4) ? din : __powf(din, 1.414f);
          v += w * dv;
        } else {
          v += 1.0f * dv;
        }
        ++slot;
      }
    }
    __stcg(out + n, v);
    const int r = receiver_of<K>(graph, n, nx, ny);
    if (r < 0) return;
    __threadfence();  // publish out[n] before the counter moves
    const int before = atomicSub(pending + r, 1);
    if (before != 1) return;  // other donors still pending: their last walker takes over
    __threadfence();  // acquire: the other donors' values are visible
    n = r;
    leaf = false;
  }
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

size_t soilb200_accumulate_ws(int nx, int ny, int edge) {
  (void)edge;
  if (nx <= 0 || ny <= 0) return 0;
  return accumulate_tiled_ws(nx, ny, false);  // 12 B/cell; the global walk alone needs 4 B/cell
}

size_t soilb200_accumulate_decay_ws(int nx, int ny, int edge) {
  (void)edge;
  if (nx <= 0 || ny <= 0) return 0;
  return accumulate_tiled_ws(nx, ny, true);  // 20 B/cell
}

int soilb200_accumulate_f32(const int32_t* graph, const float* source, const float* decay, float* out, int nx,
                            int ny, int edge, void* ws, size_t ws_bytes, void* stream) {
  if (!graph || !source || !out || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if (edge != SOILB200_D4 && edge != SOILB200_D8) return SOILB200_E_EDGE;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  if (!ws || ws_bytes < (size_t)nx * ny * sizeof(int)) return SOILB200_E_WORKSPACE;
  cudaStream_t s = (cudaStream_t)stream;
  // with the full workspace: shared-memory tiles (accumulate_tiled.cu); with only the 4 B/cell counters the global
  // walk below.  (Maps narrower than three columns always walk: the tiles tell neighbours apart by their flat offsets,
  // and with ny < 3 a diagonal offset equals a row offset.)
  if (ny >= 3 && ws_bytes >= accumulate_tiled_ws(nx, ny, decay != nullptr))
    return launch_accumulate_tiled(graph, source, decay, out, nx, ny, edge, ws, s);
  int* pending = (int*)ws;
  const long long N = (long long)nx * ny;
  const dim3 grid((ny + 31) / 32, (nx + 7) / 8);
  if (edge == SOILB200_D4) k_indegree<4><<<grid, 256, 0, s>>>(graph, source, pending, out, nx, ny);
  else k_indegree<8><<<grid, 256, 0, s>>>(graph, source, pending, out, nx, ny);
  const unsigned nb = blocks_for(N, 256);
  if (edge == SOILB200_D4) {
    if (decay) k_walk<4, true><<<nb, 256, 0, s>>>(graph, source, decay, pending, out, nx, ny);
    else k_walk<4, false><<<nb, 256, 0, s>>>(graph, source, nullptr, pending, out, nx, ny);
  } else {
    if (decay) k_walk<8, true><<<nb, 256, 0, s>>>(graph, source, decay, pending, out, nx, ny);
    else k_walk<8, false><<<nb, 256, 0, s>>>(graph, source, nullptr, pending, out, nx, ny);
  }
  count_launch(2);
  return check_launch();
}

// ---- where a cell drains to --------------------------------------------------------------------------
// Strip-partitioned accumulation (soillib_b200/accumulate_strips.py) needs, for the few cells on a strip's
// boundary rows, the last cell of their flow path inside the strip: one thread per start cell follows the
// receivers until it meets a cell without one.  A path is a chain of dependent loads, but there are only
// ~2 * ny of them; the step cap keeps a cyclic (invalid) graph from hanging the GPU.
__global__ void __launch_bounds__(128) k_terminal(const int32_t* __restrict__ graph, const int32_t* __restrict__ starts,
                                                  int32_t* __restrict__ out, int n_starts, long long n_cells, int nx,
                                                  int ny) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_starts) return;
  int c = __ldg(starts + i);
  if (c < 0 || c >= n_cells) {
    out[i] = -1;
    return;
  }
  for (long long step = 0; step < n_cells; ++step) {
    // only registered edges are followed, as in accumulate: the receiver must be one of the 8 neighbours
    const int r = receiver_of<8>(graph, c, nx, ny);
    if (r < 0) break;
    c = r;
  }
  out[i] = c;
}

// ---- accumulate without any allocation ------------------------------------------------------------
//
// For callers that are not allowed to obtain device memory at all: a host application that hands this
// library a fixed budget and forbids cudaMalloc in its processing loop (cudaMalloc takes a device-wide
// lock and synchronises, which a frame- or latency-bound host cannot afford), or one that embeds the
// library where no allocator is reachable (no torch, no pool, just pointers to its own input and output
// maps).  The pending-donor counters that soilb200_accumulate_f32 takes as caller workspace live here in
// a fixed array inside the library image instead: 4 bytes x SOILB200_ACCUMULATE_STATIC_MAX_CELLS = 16 MiB
// of device memory, held by every GPU that loads the library until the process exits.
static __device__ __align__(256) int g_static_pending[SOILB200_ACCUMULATE_STATIC_MAX_CELLS];

// Device address of g_static_pending per device ordinal.  In host code the array's name is only a handle
// for the runtime, not an address a kernel could use; cudaGetSymbolAddress translates it for the device
// that is current at the time of the call.
static int* g_static_pending_addr[64] = {};

int soilb200_graph_terminal_i32(const int32_t* graph, const int32_t* starts, int32_t* out, int n_starts, int nx, int ny,
                                void* stream) {
  if (!graph || !starts || !out || n_starts < 0 || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  if (n_starts == 0) return SOILB200_OK;
  k_terminal<<<blocks_for(n_starts, 128), 128, 0, (cudaStream_t)stream>>>(graph, starts, out, n_starts, (long long)nx * ny, nx, ny);
  count_launch();
  return check_launch();
}

int soilb200_accumulate_static_f32(const int32_t* graph, const float* source, const float* decay, float* out,
                                   int nx, int ny, int edge, void* stream) {
  if (!graph || !source || !out || nx <= 0 || ny <= 0) return SOILB200_E_ARG;
  if (edge != SOILB200_D4 && edge != SOILB200_D8) return SOILB200_E_EDGE;
  if ((long long)nx * ny > (long long)SOILB200_ACCUMULATE_STATIC_MAX_CELLS) return SOILB200_E_LIMIT;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return SOILB200_E_LIMIT;
  if (!g_static_pending_addr[dev]) {  // first call on this device: resolve the array's device address once
    int* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_static_pending);
    if (e != cudaSuccess) return (int)e;
    g_static_pending_addr[dev] = p;
  }
  // Same two kernels as soilb200_accumulate_f32.  k_indegree writes pending[n] for every n < nx*ny before
  // k_walk reads it, so nothing left in the array by an earlier call (of any size) is ever observed.
  return soilb200_accumulate_f32(graph, source, decay, out, nx, ny, edge, g_static_pending_addr[dev],
                                 (size_t)nx * ny * sizeof(int), stream);
}

}  // extern "C"
