#pragma once
#include <silt/core/tensor.hpp>
namespace silt {
template <typename T> __global__ void shim_fill(T* p, const T v, const size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
template <typename T, typename U> void set(tensor_t<T> t, const U value) {
  const size_t n = t.elem();
  shim_fill<T><<<(unsigned)((n + 511) / 512), 512>>>(t.data(), (T)value, n);
}
template <typename T> void set(tensor_t<T> dst, const tensor_t<T> src) {
  cudaMemcpyAsync(dst.data(), src.data(), src.elem() * sizeof(T), cudaMemcpyDeviceToDevice, 0);
}
}  // namespace silt
