// multiflow.cu - the multi-flow accumulation loop of example/dem_multiflow.py:43-49 kept on the device.
//
// The reference script draws K Gibbs-random receiver graphs, accumulates each, copies every result to the
// host and averages there (K D2H copies of the whole map, K curand_init passes with per-cell skip-ahead,
// K x 2N(8+8K) bytes of scratch churn).  Here one call enqueues, per sample k: random_weighted (continuing
// the per-cell XORWOW streams kept in the workspace), accumulate, and out += double(acc / K_total) - the
// script's arithmetic (`multiflow += accumulation.numpy() / float(K)`: binary32 quotient, binary64 running
// sum, k ascending), so the mean is bit-identical to the script's when one call covers all K samples.
// A range [k_first, k_first + k_count) lets replicas on several GPUs split the samples and all-reduce `out`.
#include "common.cuh"

namespace sb {
namespace {

__global__ void __launch_bounds__(256) k_mean_add(const float* __restrict__ acc, double* __restrict__ out, long long n,
                                                  float k_total, bool first) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double term = (double)(__ldcs(acc + i) / k_total);
  out[i] = first ? term : out[i] + term;
}

inline size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" {

size_t soilb200_multiflow_ws(int nx, int ny, int edge) {
  if (nx <= 0 || ny <= 0) return 0;
  const size_t n = (size_t)nx * ny;
  return align256(n * sizeof(int32_t)) + align256(n * sizeof(float)) + align256(soilb200_random_weighted_ws(nx, ny)) +
         align256(soilb200_accumulate_ws(nx, ny, edge));
}

int soilb200_multiflow_f64(const float* height, const float* source, double* out, int nx, int ny, int edge,
                           uint64_t seed, uint64_t k_first, uint64_t k_count, uint64_t k_total, float temperature,
                           int add_to_out, void* ws, size_t ws_bytes, void* stream) {
  if (!height || !source || !out || nx <= 0 || ny <= 0 || k_total == 0) return SOILB200_E_ARG;
  if (edge != SOILB200_D4 && edge != SOILB200_D8) return SOILB200_E_EDGE;
  if ((long long)nx * ny > 0x7fffffffLL) return SOILB200_E_LIMIT;
  if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255) || ws_bytes < soilb200_multiflow_ws(nx, ny, edge))
    return SOILB200_E_WORKSPACE;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t n = (size_t)nx * ny;
  char* p = (char*)ws;
  int32_t* graph = (int32_t*)p;       p += align256(n * sizeof(int32_t));
  float* acc = (float*)p;             p += align256(n * sizeof(float));
  void* rw_ws = p;                    const size_t rw_bytes = soilb200_random_weighted_ws(nx, ny); p += align256(rw_bytes);
  void* acc_ws = p;                   const size_t acc_bytes = soilb200_accumulate_ws(nx, ny, edge);
  // a fresh header: the first sample seeds every stream at offset k_first, the following ones continue it
  cudaError_t e = cudaMemsetAsync(rw_ws, 0, 64, s);
  if (e != cudaSuccess) return (int)e;
  if (k_count == 0 && !add_to_out) {
    e = cudaMemsetAsync(out, 0, n * sizeof(double), s);
    return e == cudaSuccess ? SOILB200_OK : (int)e;
  }
  for (uint64_t k = k_first; k < k_first + k_count; ++k) {
    int rc = soilb200_random_weighted_f32(height, graph, nx, ny, edge, seed, k, temperature, rw_ws, rw_bytes, s);
    if (rc != SOILB200_OK) return rc;
    rc = soilb200_accumulate_f32(graph, source, nullptr, acc, nx, ny, edge, acc_ws, acc_bytes, s);
    if (rc != SOILB200_OK) return rc;
    k_mean_add<<<blocks_for((long long)n, 256), 256, 0, s>>>(acc, out, (long long)n, (float)k_total,
                                                             k == k_first && !add_to_out);
    count_launch();
  }
  return check_launch();
}

}  // extern "C"
