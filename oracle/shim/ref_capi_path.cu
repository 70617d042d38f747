// oracle/shim/ref_capi_path.cu - TEST INFRASTRUCTURE ONLY.
// Third translation unit of the reference shim: path.cu (soil::solve_uniform), compiled UNMODIFIED from
// /root/reference behind a C API.  It defines its own __stepsize and block helpers, hence its own TU.
#include <silt/silt.hpp>
#include <soillib/model/path/path.cu>

extern "C" {

// flow (nx,ny,2), source (nx,ny,K), decay (nx,ny), rng: n_samples curandState; flux_out (nx,ny,K)
int ref_solve_uniform(const float* flow, const float* source, const float* decay, void* rng, long n_samples,
                      float* flux_out, int nx, int ny, int K, const float* scale, long count) {
  auto r = soil::solve_uniform(silt::tensor_t<float>((float*)flow, silt::shape(nx, ny, 2)),
                               silt::tensor_t<float>((float*)source, silt::shape(nx, ny, K)),
                               silt::tensor_t<float>((float*)decay, silt::shape(nx, ny)),
                               silt::tensor_t<silt::rng>((silt::rng*)rng, silt::shape(n_samples)),
                               silt::vec2(scale[0], scale[1]), (size_t)count);
  cudaMemcpy(flux_out, r.as<float>().data(), sizeof(float) * K * (size_t)nx * ny, cudaMemcpyDeviceToDevice);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}

}  // extern "C"
