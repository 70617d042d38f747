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

// Division by a divisor that is the same for every cell of a launch (the cell scale, the normalisation length).
// The IEEE division costs a reciprocal approximation, two Newton steps, the quotient, its residual and a correction,
// plus a range check with a slow path: ~10 instructions, and a dozen of them per cell make the fused step issue-bound.
// With the divisor known on the host:
//   kind 1  the divisor is a power of two: x * (1/d) is exact, hence identical to x / d for every x;
//   kind 2  r = RN(1/d) computed once on the host, then q0 = x * r, rem = fma(-q0, d, x), q = fma(rem, r, q0) - the tail
//           of the sequence the compiler emits for a division (which refines its reciprocal to the same r): the
//           correctly rounded quotient whenever nothing under- or overflows on the way, which the guard on |x| and
//           the limits on |d| make sure of; numerators outside the guard (zeros, denormals, huge values, NaN) take
//           the plain division.  soilb200_selftest_div_const runs all 2^32 numerators through both for a divisor;
//   kind 0  anything else (and every scale3 not built by make_scale): the plain division.
struct scale3 {
  float x, y, z;
  float rx = 0.0f, ry = 0.0f, rz = 0.0f;  // reciprocals, as far as kind != 0
  int kx = 0, ky = 0, kz = 0;
};

inline int divisor_kind(float d, float* r) {
  *r = 0.0f;
  if (!(d == d) || d == 0.0f) return 0;
  const float a = d < 0.0f ? -d : d;
  if (!(a >= 9.5367431640625e-07f && a <= 1048576.0f)) return 0;  // 2^-20 .. 2^20
  *r = 1.0f / d;                                                  // correctly rounded (host code, no contraction)
  unsigned bits;
  static_assert(sizeof(bits) == sizeof(a), "binary32");
  __builtin_memcpy(&bits, &a, sizeof(bits));
  return (bits & 0x007fffffu) == 0u ? 1 : 2;
}
inline scale3 make_scale(float x, float y, float z, bool fast = true) {
  scale3 s{x, y, z};
  if (fast) {
    s.kx = divisor_kind(x, &s.rx);
    s.ky = divisor_kind(y, &s.ry);
    s.kz = divisor_kind(z, &s.rz);
  }
  return s;
}

#ifdef __CUDACC__
// x / d for a launch-constant divisor d (see above); identical to the IEEE quotient bit for bit
__device__ __forceinline__ float cdiv(float x, float d, float r, int kind) {
  if (kind == 1) return x * r;
  if (kind == 2) {
    const float ax = fabsf(x);
    if (ax >= 1.0e-28f && ax <= 1.0e30f) {
      const float q0 = x * r;
      const float rem = __fmaf_rn(-q0, d, x);
      return __fmaf_rn(rem, r, q0);
    }
    if (x == 0.0f) return d > 0.0f ? x : -x;  // a signed zero, no slow path
  }
  return x / d;
}
#endif

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
