// oracle/shim/ref_capi_grad.cu - TEST INFRASTRUCTURE ONLY.
// Second translation unit of the reference shim: grad.cu and filter.cu (each defines the same file-local helper as
// graph.cu, so they cannot share a TU with it), compiled UNMODIFIED from /root/reference behind a C API.
#include <silt/silt.hpp>
#include <soillib/model/grad/grad.cu>
#define block block_filter  // filter.cu repeats grad.cu's anonymous-namespace helper under the same name
#include <soillib/model/filter/filter.cu>
#undef block

namespace {
template <typename T> silt::tensor_t<T> wrap2(const void* p, size_t a, size_t b) { return silt::tensor_t<T>((T*)p, silt::shape(a, b)); }
template <typename T> silt::tensor_t<T> wrap3(const void* p, size_t a, size_t b, size_t c) { return silt::tensor_t<T>((T*)p, silt::shape(a, b, c)); }
}

extern "C" {

int ref_gradient(const float* h, float* out, int nx, int ny, const float* scale) {
  auto r = soil::gradient(wrap2<float>(h, nx, ny), silt::vec2(scale[0], scale[1]));
  cudaMemcpyAsync(out, r.data(), sizeof(float) * 2 * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_negslope(const float* h, float* out, int nx, int ny, const float* scale) {
  auto r = soil::negslope(wrap2<float>(h, nx, ny), silt::vec2(scale[0], scale[1]));
  cudaMemcpyAsync(out, r.data(), sizeof(float) * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_laplacian(const float* in, float* out, int nx, int ny, int channels, const float* scale) {
  auto r = soil::laplacian(wrap3<float>(in, nx, ny, channels), silt::vec2(scale[0], scale[1]));
  cudaMemcpyAsync(out, r.data(), sizeof(float) * channels * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
// gaussian_blur overwrites AND returns its argument (filter.cu:82,90): `inout` holds the result afterwards
int ref_gaussian_blur(float* inout, int nx, int ny, int channels, float sigma) {
  soil::gaussian_blur(wrap3<float>(inout, nx, ny, channels), sigma);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}

}  // extern "C"
