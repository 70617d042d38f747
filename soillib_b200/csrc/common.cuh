// common.cuh - conventions shared by every kernel of the hot path.
//
// Layout: row-major, flatten(x,y) = x*ny + y (dim 1 fastest),
// trailing vector dimension interleaved.  All arithmetic is IEEE binary32 "as
// written in the reference": the library is compiled with -fmad=false so that
// no multiply-add is contracted and results are reproducible against the CPU
// oracle; the reference's fast intrinsics (__expf, __powf) are kept as such.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>
#include "../../include/soilb200.h"

namespace sb {

struct grid2 {
  int nx, ny;
  __host__ __device__ __forceinline__ bool oob(int x, int y) const { return x < 0 || y < 0 || x >= nx || y >= ny; }
  __host__ __device__ __forceinline__ long long elem() const { return (long long)nx * ny; }
};

struct scale3 {
  float x, y, z;
};

// neighbour table, graph.hpp:21-46
__device__ __constant__ const int c_shift_x[8] = {-1, 0, 0, 1, -1, -1, 1, 1};
__device__ __constant__ const int c_shift_y[8] = {0, -1, 1, 0, -1, 1, -1, 1};

extern unsigned long long g_launches;  // counted on the host, one per kernel launch
inline void count_launch(int n = 1) { g_launches += (unsigned long long)n; }

inline int check_launch() {
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? SOILB200_OK : (int)e;
}

inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

inline int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

// what the fused step knows about the row strip it runs on (multi-GPU); see transport.cu strip_t.
// Whole domain: row0 = 0, own = [0, nx), nx_g = nx, no outboxes.
struct strip_host {
  int row0, own_lo, own_hi, nx_g;
  float *outbox_lo, *outbox_hi;
  int* counts;
  int capacity;
  const void* peers;  // device array of n_peers soilb200_peer_t (every rank's maps, peer-mapped), or null
  int n_peers, peer_self;
};

// sqrt(x*x + y*y), erosion_map.cu:49-53 / graph.cu:23-25
__device__ __forceinline__ float len2(float x, float y) { return sqrtf(x * x + y * y); }

// streaming accessors: data touched once per pass should not displace the
// stencil rows in L1
__device__ __forceinline__ float ld_stream(const float* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(float* p, float v) { __stcs(p, v); }

}  // namespace sb
