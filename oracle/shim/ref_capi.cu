// oracle/shim/ref_capi.cu - TEST INFRASTRUCTURE ONLY.
// Compiles the reference's own translation units UNMODIFIED, from where they
// lie under /root/reference, and exposes them behind a C API that takes raw
// device pointers.  Output goes to oracle/_ref/ (git-ignored).  Used as
//   (1) the GPU-side oracle (same intrinsics as the reference => tight parity),
//   (2) the "reference's own kernels on the same B200" baseline.
#include <silt/silt.hpp>
#include <soillib/model/graph/graph.cu>
#include <soillib/model/path/erosion.cu>
#include <cstring>

namespace {
template <typename T> silt::tensor_t<T> wrap(const void* p, size_t a, size_t b) { return silt::tensor_t<T>((T*)p, silt::shape(a, b)); }
template <typename T> silt::tensor_t<T> wrap(const void* p, size_t a, size_t b, size_t c) { return silt::tensor_t<T>((T*)p, silt::shape(a, b, c)); }
template <typename T> silt::tensor_t<T> wrap1(const void* p, size_t a) { return silt::tensor_t<T>((T*)p, silt::shape(a)); }
soil::param_t to_param(const void* p) { soil::param_t q; static_assert(sizeof(soil::param_t) == 112, "param_t"); std::memcpy(&q, p, sizeof(q)); return q; }
__global__ void k_add(float* a, const float* b, size_t n) { size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; if (i < n) a[i] += b[i]; }
}

extern "C" {

int ref_direction(const float* h, int* out, int nx, int ny, int edge) {
  auto r = soil::direction(wrap<float>(h, nx, ny), (soil::edge_t)edge);
  cudaMemcpyAsync(out, r.data(), sizeof(int) * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_steepest(const float* h, int* out, int nx, int ny, int edge) {
  auto r = soil::steepest(wrap<float>(h, nx, ny), (soil::edge_t)edge);
  cudaMemcpyAsync(out, r.data(), sizeof(int) * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_random_weighted(const float* h, int* out, int nx, int ny, int edge, unsigned long long seed, unsigned long long offset, float T) {
  auto r = soil::random_weighted(wrap<float>(h, nx, ny), (soil::edge_t)edge, seed, offset, T);
  cudaMemcpyAsync(out, r.data(), sizeof(int) * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_accumulate(const int* graph, const float* src, const float* decay, float* out, int nx, int ny, int edge) {
  auto r = decay ? soil::accumulate_decay(wrap<int>(graph, nx, ny), wrap<float>(src, nx, ny), wrap<float>(decay, nx, ny), (soil::edge_t)edge)
                 : soil::accumulate(wrap<int>(graph, nx, ny), wrap<float>(src, nx, ny), (soil::edge_t)edge);
  cudaMemcpyAsync(out, r.data(), sizeof(float) * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_slope(const float* h, const int* flow, float* out, int nx, int ny, const float* scale) {
  auto r = soil::slope(wrap<float>(h, nx, ny), wrap<int>(flow, nx, ny), silt::vec2(scale[0], scale[1]));
  cudaMemcpyAsync(out, r.data(), sizeof(float) * (size_t)nx * ny, cudaMemcpyDeviceToDevice, 0);
  silt::pool::recycle();
  return (int)cudaGetLastError();
}
int ref_seed(void* rng, long n, unsigned long long seed, unsigned long long offset) {
  soil::__seed<<<(unsigned)((n + 511) / 512), 512>>>(wrap1<curandState>(rng, n), seed, offset);
  return (int)cudaGetLastError();
}

// argument order = erosion.hpp:69-84
int ref_transport_fluvial(float* layers, float* rainfall, float* discharge, float* discharge_track, float* mass, float* mass_track,
                          float* momentum, float* momentum_track, float* albedo_bedrock, float* albedo_transport, float* albedo_surface,
                          void* rng, long P, int nx, int ny, const float* scale, const void* param) {
  soil::transport_fluvial(wrap<float>(layers, nx, ny, 2), wrap<float>(rainfall, nx, ny), wrap<float>(discharge, nx, ny),
                          wrap<float>(discharge_track, nx, ny), wrap<float>(mass, nx, ny), wrap<float>(mass_track, nx, ny),
                          wrap<float>(momentum, nx, ny, 2), wrap<float>(momentum_track, nx, ny, 2), wrap<float>(albedo_bedrock, nx, ny, 3),
                          wrap<float>(albedo_transport, nx, ny, 3), wrap<float>(albedo_surface, nx, ny, 3), wrap1<silt::rng>(rng, P),
                          silt::vec3(scale[0], scale[1], scale[2]), to_param(param));
  return (int)cudaGetLastError();
}
// erosion.hpp:86-98
int ref_transport_debris(float* layers, float* velocity, float* velocity_track, float* mass, float* mass_track, float* albedo_bedrock,
                         float* albedo_transport, float* albedo_surface, void* rng, long P, int nx, int ny, const float* scale, const void* param) {
  soil::transport_debris(wrap<float>(layers, nx, ny, 2), wrap<float>(velocity, nx, ny, 2), wrap<float>(velocity_track, nx, ny, 2),
                         wrap<float>(mass, nx, ny), wrap<float>(mass_track, nx, ny), wrap<float>(albedo_bedrock, nx, ny, 3),
                         wrap<float>(albedo_transport, nx, ny, 3), wrap<float>(albedo_surface, nx, ny, 3), wrap1<silt::rng>(rng, P),
                         silt::vec3(scale[0], scale[1], scale[2]), to_param(param));
  return (int)cudaGetLastError();
}
// erosion.hpp:104-119
int ref_mass_transfer(float* delta, float* layers, const float* uplift, const float* discharge, const float* mass, const float* momentum,
                      const float* debris, const float* momentumDebris, float* albedo_bedrock, float* albedo_fluvial, float* albedo_debris,
                      float* albedo_surface, int nx, int ny, const float* scale, const void* param) {
  soil::mass_transfer(wrap<float>(delta, nx, ny, 2), wrap<float>(layers, nx, ny, 2), wrap<float>(uplift, nx, ny), wrap<float>(discharge, nx, ny),
                      wrap<float>(mass, nx, ny), wrap<float>(momentum, nx, ny, 2), wrap<float>(debris, nx, ny), wrap<float>(momentumDebris, nx, ny, 2),
                      wrap<float>(albedo_bedrock, nx, ny, 3), wrap<float>(albedo_fluvial, nx, ny, 3), wrap<float>(albedo_debris, nx, ny, 3),
                      wrap<float>(albedo_surface, nx, ny, 3), silt::vec3(scale[0], scale[1], scale[2]), to_param(param));
  return (int)cudaGetLastError();
}
int ref_mass_creep(float* delta, const float* layers, int nx, int ny, const float* scale, const void* param) {
  soil::mass_creep(wrap<float>(delta, nx, ny, 2), wrap<float>(layers, nx, ny, 2), silt::vec3(scale[0], scale[1], scale[2]), to_param(param));
  return (int)cudaGetLastError();
}
int ref_layer_merge(float* height, const float* layers, int nx, int ny) {
  soil::layer_merge(wrap<float>(height, nx, ny), wrap<float>(layers, nx, ny, 2));
  return (int)cudaGetLastError();
}

// One erosion step in the reference's unfused sequence.
// Field order matches oracle/soil_oracle.cu model_t.
struct ref_model_t {
  int nx, ny;
  float scale[3];
  float *layers, *height, *rainfall, *uplift;
  float *waterHeight, *waterFlux, *mass, *massFlux, *momentum, *velocityFlux;
  float *debris, *debrisFlux, *debrisVel, *debrisVelFlux;
  float *albedoBedrock, *albedoSurface, *albedoFluxFluvial, *albedoFluxDebris;
  float* delta;
  void *rngFluvial, *rngDebris;
  long P;
};

int ref_erode_step(ref_model_t* m, const void* param, int with_creep) {
  const size_t N = (size_t)m->nx * m->ny;
  cudaMemsetAsync(m->waterFlux, 0, N * 4, 0);
  cudaMemsetAsync(m->massFlux, 0, N * 4, 0);
  cudaMemsetAsync(m->velocityFlux, 0, N * 8, 0);
  cudaMemsetAsync(m->debrisFlux, 0, N * 4, 0);
  cudaMemsetAsync(m->debrisVelFlux, 0, N * 8, 0);
  cudaMemsetAsync(m->albedoFluxFluvial, 0, N * 12, 0);
  cudaMemsetAsync(m->albedoFluxDebris, 0, N * 12, 0);
  ref_transport_fluvial(m->layers, m->rainfall, m->waterHeight, m->waterFlux, m->mass, m->massFlux, m->momentum, m->velocityFlux,
                        m->albedoBedrock, m->albedoFluxFluvial, m->albedoSurface, m->rngFluvial, m->P, m->nx, m->ny, m->scale, param);
  ref_transport_debris(m->layers, m->debrisVel, m->debrisVelFlux, m->debris, m->debrisFlux, m->albedoBedrock, m->albedoFluxDebris,
                       m->albedoSurface, m->rngDebris, m->P, m->nx, m->ny, m->scale, param);
  cudaMemsetAsync(m->delta, 0, N * 8, 0);
  ref_mass_transfer(m->delta, m->layers, m->uplift, m->waterHeight, m->mass, m->momentum, m->debris, m->debrisVel, m->albedoBedrock,
                    m->albedoFluxFluvial, m->albedoFluxDebris, m->albedoSurface, m->nx, m->ny, m->scale, param);
  if (with_creep) ref_mass_creep(m->delta, m->layers, m->nx, m->ny, m->scale, param);
  k_add<<<(unsigned)((2 * N + 511) / 512), 512>>>(m->layers, m->delta, 2 * N);
  ref_layer_merge(m->height, m->layers, m->nx, m->ny);
  return (int)cudaGetLastError();
}

int ref_sync() { return (int)cudaDeviceSynchronize(); }
void ref_drain_pool() { silt::pool::drain(); }


// erosion.cu:828-919
int ref_albedo_layer(float* albedo, const float* bedrock, const float* sediment, const float* layers, int nx, int ny,
                     float scaleSediment, const float* shift) {
  soil::albedo_layer(wrap<float>(albedo, nx, ny, 3), wrap<float>(bedrock, nx, ny, 3), wrap<float>(sediment, nx, ny, 3),
                     wrap<float>(layers, nx, ny, 2), scaleSediment, silt::vec3(shift[0], shift[1], shift[2]));
  return (int)cudaGetLastError();
}
int ref_albedo_stratum(float* bedrock, const float* uplift, const float* layers, int nx, int ny, const float* scale,
                       const void* param, const float* a, const float* b, float age, float freq) {
  soil::albedo_stratum(wrap<float>(bedrock, nx, ny, 3), wrap<float>(uplift, nx, ny), wrap<float>(layers, nx, ny, 2),
                       silt::vec3(scale[0], scale[1], scale[2]), to_param(param), silt::vec3(a[0], a[1], a[2]),
                       silt::vec3(b[0], b[1], b[2]), age, freq);
  return (int)cudaGetLastError();
}
int ref_albedo_discharge(float* albedo, const float* discharge, int nx, int ny, const float* col, float extinction,
                         float scale) {
  soil::albedo_discharge(wrap<float>(albedo, nx, ny, 3), wrap<float>(discharge, nx, ny),
                         silt::vec3(col[0], col[1], col[2]), extinction, scale);
  return (int)cudaGetLastError();
}

}  // extern "C"
