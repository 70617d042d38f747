// micro_gather.cu - ceilings for the particle kernel's access pattern on one GPU (no library code involved).
//
// A "walker" is a dependent chain of crossings on an n x n map: every crossing gathers what a particle gathers
// (5-point height stencil + velocity + water height of the cell it stands on), optionally posts the two 16-byte
// reductions + tag bit a fluvial particle posts, and moves to a 4-neighbour chosen from the loaded bits (so the
// next address depends on this crossing's data, like a particle's).  Layouts compared:
//   scalar7  the library's round-1 layout: row-major height, discharge, momentum maps, 7 scalar loads
//   tile16   one 16-byte record {h, wh, vx, vy} per cell in 8x8-cell tiles (1 KB): 1 x LDG.128 + 4 x LDG.32
//   rec32    one 32-byte record {h, h(x-1), h(x+1), h(y-1), h(y+1), wh, vx, vy} per cell: 2 x LDG.128
//   blk2x2   heights in 2x2-cell blocks (16 B) + {wh, vx, vy, -}: 4 x LDG.128
//   morton16 the 16-byte records in Morton (Z) order inside 8x8-cell tiles: a 2x2-cell block is one 64-byte DRAM
//            access, a 2x4 block one 128-byte line, a 4x4 block one 256-byte interleave chunk
// Second part: read throughput of random aligned chunks of 32 B ... 2 KB (what a burst of a tiled layout is worth).
// Reported: crossings/s for T walkers (T = 64 Ki is the particle kernel's thread count at P = 32768 x 2 systems).
// If the rate grows with T the chain is latency-bound; where it saturates is the throughput ceiling of the layout.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/micro_gather scripts/micro_gather.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

static const int n = 8192;

__device__ __forceinline__ unsigned mix(unsigned a) {
  a ^= a >> 16; a *= 0x7feb352dU; a ^= a >> 15; a *= 0x846ca68bU; a ^= a >> 16;
  return a;
}

struct maps {
  const float* h; const float* wh; const float2* v;      // scalar7
  const float4* t16;                                      // tile16
  const float4* r32;                                      // rec32 (2 float4 per cell)
  const float4* b22; const float4* c16;                   // blk2x2 + per-cell {wh,vx,vy,-}
  float* flux; unsigned* tags;                            // 64 B per cell, 1 bit per cell
};

__device__ __forceinline__ size_t tile_index(int x, int y) {  // 8x8 tiles, row-major inside
  return ((size_t)(x >> 3) * (n >> 3) + (y >> 3)) * 64 + ((x & 7) << 3) + (y & 7);
}
__device__ __forceinline__ size_t morton_index(int x, int y) {  // 8x8 tiles row-major, Z order inside
  const unsigned a = x & 7, b = y & 7;
  const unsigned z = (b & 1) | ((a & 1) << 1) | ((b & 2) << 1) | ((a & 2) << 2) | ((b & 4) << 2) | ((a & 4) << 3);
  return ((size_t)(x >> 3) * (n >> 3) + (y >> 3)) * 64 + z;
}
__device__ __forceinline__ float h_blk(const float4* b22, int x, int y) {
  const float4 q = __ldg(b22 + (size_t)(x >> 1) * (n >> 1) + (y >> 1));
  const int k = ((x & 1) << 1) | (y & 1);
  return k == 0 ? q.x : k == 1 ? q.y : k == 2 ? q.z : q.w;
}

template <int LAYOUT, bool RED>
__global__ void __launch_bounds__(64) k_walk(maps m, int T, int iters, float* sink) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  unsigned s = mix(t * 2654435761u + 12345u);
  int x = 2 + (s % (n - 4)), y = 2 + ((s >> 13) % (n - 4));
  int dir = s & 3;
  float acc = 0.0f;
  for (int it = 0; it < iters; ++it) {
    float c, xm, xp, ym, yp, w, vx, vy;
    if (LAYOUT == 0) {
      const size_t i = (size_t)x * n + y;
      c = __ldg(m.h + i); xm = __ldg(m.h + i - n); xp = __ldg(m.h + i + n); ym = __ldg(m.h + i - 1); yp = __ldg(m.h + i + 1);
      const float2 v = __ldg(m.v + i); vx = v.x; vy = v.y; w = __ldg(m.wh + i);
    } else if (LAYOUT == 1) {
      const float4 r = __ldg(m.t16 + tile_index(x, y));
      c = r.x; w = r.y; vx = r.z; vy = r.w;
      xm = __ldg(reinterpret_cast<const float*>(m.t16 + tile_index(x - 1, y)));
      xp = __ldg(reinterpret_cast<const float*>(m.t16 + tile_index(x + 1, y)));
      ym = __ldg(reinterpret_cast<const float*>(m.t16 + tile_index(x, y - 1)));
      yp = __ldg(reinterpret_cast<const float*>(m.t16 + tile_index(x, y + 1)));
    } else if (LAYOUT == 2) {
      const size_t i = (size_t)x * n + y;
      const float4 a = __ldg(m.r32 + 2 * i), b = __ldg(m.r32 + 2 * i + 1);
      c = a.x; xm = a.y; xp = a.z; ym = a.w; yp = b.x; w = b.y; vx = b.z; vy = b.w;
    } else if (LAYOUT == 4) {
      const float4 r = __ldg(m.t16 + morton_index(x, y));
      c = r.x; w = r.y; vx = r.z; vy = r.w;
      xm = __ldg(reinterpret_cast<const float*>(m.t16 + morton_index(x - 1, y)));
      xp = __ldg(reinterpret_cast<const float*>(m.t16 + morton_index(x + 1, y)));
      ym = __ldg(reinterpret_cast<const float*>(m.t16 + morton_index(x, y - 1)));
      yp = __ldg(reinterpret_cast<const float*>(m.t16 + morton_index(x, y + 1)));
    } else if (LAYOUT == 5) {  // no gathers at all: what the reductions alone cost
      c = xm = xp = ym = yp = w = vx = vy = __uint_as_float(s & 0x3fffffffu);
    } else {
      c = h_blk(m.b22, x, y);
      // the block of (x,y) holds one x-neighbour and one y-neighbour; two more blocks hold the others
      xm = h_blk(m.b22, x - 1, y); xp = h_blk(m.b22, x + 1, y); ym = h_blk(m.b22, x, y - 1); yp = h_blk(m.b22, x, y + 1);
      const float4 r = __ldg(m.c16 + (size_t)x * n + y);
      w = r.x; vx = r.y; vy = r.z;
    }
    const float g = (c - xm) + (xp - c) * 0.5f + (c - ym) * 0.25f + (yp - c) * 0.125f + w + vx + vy;
    acc += g;
    if (RED) {
      const size_t i = (size_t)x * n + y;
      float* r = m.flux + 16 * i;
      atomicAdd(reinterpret_cast<float4*>(r), make_float4(g, 1.0f, vx, vy));
      atomicAdd(reinterpret_cast<float4*>(r + 4), make_float4(w, c, 1.0f, 0.0f));
      atomicOr(m.tags + (i >> 5), 1u << (i & 31));
    }
    // next cell: keep the direction 7 times out of 8, the choice depends on the loaded bits
    s = mix(s + __float_as_uint(g));
    if ((s & 7) == 0) dir = (s >> 3) & 3;
    x += dir == 0 ? 1 : dir == 1 ? -1 : 0;
    y += dir == 2 ? 1 : dir == 3 ? -1 : 0;
    if (x < 2) { x = 2; dir = 0; } if (x > n - 3) { x = n - 3; dir = 1; }
    if (y < 2) { y = 2; dir = 2; } if (y > n - 3) { y = n - 3; dir = 3; }
  }
  if (acc == 12345.678f) sink[0] = acc;
}

template <int LAYOUT, bool RED>
static void run(const char* name, const maps& m, int T, int iters, float* sink) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_walk<LAYOUT, RED><<<(T + 63) / 64, 64>>>(m, T, iters, sink);
  cudaEventRecord(e0);
  k_walk<LAYOUT, RED><<<(T + 63) / 64, 64>>>(m, T, iters, sink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("%-8s red=%d T=%8d iters=%4d  %8.3f ms  %7.2f G crossings/s  %7.1f ns/crossing/walker  (%s)\n", name, (int)RED, T, iters,
         ms, (double)T * iters / ms / 1e6, ms * 1e6 / iters, cudaGetErrorString(cudaGetLastError()));
}

// random aligned chunks: every warp reads `per_warp` chunks of `chunk` bytes, 16 bytes per lane and load
__global__ void __launch_bounds__(256) k_chunks(const float4* buf, size_t n_chunks, int chunk, int per_warp, float* sink) {
  const int lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lanes = chunk / 16 < 32 ? chunk / 16 : 32;      // lanes that share one chunk
  const int groups = 32 / lanes;                            // chunks a warp reads per load instruction
  const int loads = chunk / 16 / lanes;                     // loads per lane and chunk (chunk > 512 B)
  float acc = 0.0f;
  for (int i = 0; i < per_warp; i += groups) {
    const unsigned r = mix((unsigned)(warp * 977u + i + lane / lanes) * 2654435761u + 17u);
    const float4* p = buf + (size_t)(r % n_chunks) * (chunk / 16) + (lane % lanes);
    for (int k = 0; k < loads; ++k) {
      const float4 v = __ldg(p + k * lanes);
      acc += v.x + v.y + v.z + v.w;
    }
  }
  if (acc == 12345.678f) sink[0] = acc;
}

static void run_chunks(const float4* buf, size_t bytes, int chunk, float* sink) {
  const int warps = 148 * 64, per_warp = 4096 * 32 / (chunk / 16 < 32 ? 1 : chunk / 512) / 8;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_chunks<<<warps / 8, 256>>>(buf, bytes / chunk, chunk, per_warp, sink);
  cudaEventRecord(e0);
  k_chunks<<<warps / 8, 256>>>(buf, bytes / chunk, chunk, per_warp, sink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const int lanes = chunk / 16 < 32 ? chunk / 16 : 32;
  const double n = (double)warps * ((per_warp + 32 / lanes - 1) / (32 / lanes)) * (32 / lanes);
  printf("random %4d-byte chunks: %8.3f ms  %7.2f G chunks/s  %7.1f GB/s  (%s)\n", chunk, ms, n / ms / 1e6, n * chunk / ms / 1e6,
         cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char** argv) {
  const size_t N = (size_t)n * n;
  float *h, *wh, *sink, *flux; float2* v; float4 *t16, *r32, *b22, *c16; unsigned* tags;
  cudaMalloc(&h, N * 4); cudaMalloc(&wh, N * 4); cudaMalloc(&v, N * 8); cudaMalloc(&t16, N * 16); cudaMalloc(&r32, N * 32);
  cudaMalloc(&b22, N * 4); cudaMalloc(&c16, N * 16); cudaMalloc(&flux, N * 64); cudaMalloc(&tags, N / 8); cudaMalloc(&sink, 4);
  // any finite content will do; 0x3c = small positive floats
  cudaMemset(h, 0x3c, N * 4); cudaMemset(wh, 0x3c, N * 4); cudaMemset(v, 0x3c, N * 8); cudaMemset(t16, 0x3c, N * 16);
  cudaMemset(r32, 0x3c, N * 32); cudaMemset(b22, 0x3c, N * 4); cudaMemset(c16, 0x3c, N * 16); cudaMemset(flux, 0, N * 64);
  cudaMemset(tags, 0, N / 8);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("alloc failed\n"); return 1; }
  const maps m{h, wh, v, t16, r32, b22, c16, flux, tags};
  const int iters = 512;
  const int Ts[] = {16384, 65536, 262144, 1048576};
  for (int T : Ts) {
    const int it = T > 262144 ? 128 : iters;
    run<0, false>("scalar7", m, T, it, sink);
    run<1, false>("tile16", m, T, it, sink);
    run<2, false>("rec32", m, T, it, sink);
    run<3, false>("blk2x2", m, T, it, sink);
    run<4, false>("morton16", m, T, it, sink);
    run<4, true>("morton16", m, T, it, sink);
    run<5, true>("red only", m, T, it, sink);
    run<0, true>("scalar7", m, T, it, sink);
    run<1, true>("tile16", m, T, it, sink);
    run<2, true>("rec32", m, T, it, sink);
  }
  for (int chunk = 32; chunk <= 2048; chunk *= 2) run_chunks(r32, N * 32, chunk, sink);
  return 0;
}
