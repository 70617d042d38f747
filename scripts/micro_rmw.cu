// probe: read-then-zero patterns of the fused kernel's hot path, isolated
#include <cstdio>
#include <cuda_runtime.h>
#define N (4096LL * 4096LL)
__global__ void k_rmw1(float* a, float* out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float v = a[i]; a[i] = 0.f; if (v == 123.f) out[0] = v;
}
__global__ void k_rmw5(float* a, float* b, float2* c, float* d, float2* e, float* out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float v = a[i], w = b[i]; float2 x = c[i]; float y = d[i]; float2 z = e[i];
  a[i] = 0.f; b[i] = 0.f; c[i] = make_float2(0.f, 0.f); d[i] = 0.f; e[i] = make_float2(0.f, 0.f);
  if (v + w + x.x + x.y + y + z.x + z.y == 123.f) out[0] = v;
}
__global__ void k_rd5(const float* a, const float* b, const float2* c, const float* d, const float2* e, float* out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float v = a[i], w = b[i]; float2 x = c[i]; float y = d[i]; float2 z = e[i];
  if (v + w + x.x + x.y + y + z.x + z.y == 123.f) out[0] = v;
}
__global__ void k_wr5(float* a, float* b, float2* c, float* d, float2* e) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  a[i] = 0.f; b[i] = 0.f; c[i] = make_float2(0.f, 0.f); d[i] = 0.f; e[i] = make_float2(0.f, 0.f);
}
__global__ void k_rmw3(float* a, float* out) {  // stride-3 (albedo) read + zero
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float v = a[3 * i], w = a[3 * i + 1], x = a[3 * i + 2];
  a[3 * i] = 0.f; a[3 * i + 1] = 0.f; a[3 * i + 2] = 0.f;
  if (v + w + x == 123.f) out[0] = v;
}
// 2-D block like the fused kernel: 32 x 8
__global__ void k_rmw5_2d(float* a, float* b, float2* c, float* d, float2* e, float* out, int ny) {
  int y = blockIdx.x * 32 + (threadIdx.x & 31), x = blockIdx.y * 8 + (threadIdx.x >> 5);
  long long i = (long long)x * ny + y;
  float v = a[i], w = b[i]; float2 xx = c[i]; float yy = d[i]; float2 z = e[i];
  a[i] = 0.f; b[i] = 0.f; c[i] = make_float2(0.f, 0.f); d[i] = 0.f; e[i] = make_float2(0.f, 0.f);
  if (v + w + xx.x + xx.y + yy + z.x + z.y == 123.f) out[0] = v;
}
template <class F> float timeit(F f, const char* name, double bytes) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) f();
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
  printf("%-12s %.3f ms  %.0f GB/s  (%s)\n", name, ms, bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
  return ms;
}
int main() {
  float *a, *b, *d, *out, *alb; float2 *c, *e;
  cudaMalloc(&a, N * 4); cudaMalloc(&b, N * 4); cudaMalloc(&d, N * 4); cudaMalloc(&c, N * 8); cudaMalloc(&e, N * 8);
  cudaMalloc(&alb, N * 12); cudaMalloc(&out, 4);
  cudaMemset(a, 0, N * 4); cudaMemset(b, 0, N * 4); cudaMemset(d, 0, N * 4); cudaMemset(c, 0, N * 8); cudaMemset(e, 0, N * 8); cudaMemset(alb, 0, N * 12);
  unsigned nb = (unsigned)((N + 255) / 256);
  timeit([&] { k_rmw1<<<nb, 256>>>(a, out); }, "rmw1", N * 8.0);
  timeit([&] { k_rmw5<<<nb, 256>>>(a, b, c, d, e, out); }, "rmw5", N * 56.0);
  timeit([&] { k_rd5<<<nb, 256>>>(a, b, c, d, e, out); }, "rd5", N * 28.0);
  timeit([&] { k_wr5<<<nb, 256>>>(a, b, c, d, e); }, "wr5", N * 28.0);
  timeit([&] { k_rmw3<<<nb, 256>>>(alb, out); }, "rmw3", N * 24.0);
  timeit([&] { k_rmw5_2d<<<dim3(4096 / 32, 4096 / 8), 256>>>(a, b, c, d, e, out, 4096); }, "rmw5_2d", N * 56.0);
  return 0;
}
