// soilb200.hpp - C++23 host side of the hot path, above the C ABI of soilb200.h.
//
// It mirrors the reference's C++ interface for this path (same function names, argument order and
// error behaviour), so that source/soillib/python/model.cpp could bind these instead of its own
// kernels:
//   soil::direction / steepest / random_weighted / accumulate / accumulate_decay   graph.hpp:49-60
//   soil::transport_fluvial / transport_debris / mass_transfer / mass_creep / layer_merge
//                                                                                erosion.hpp:69-142
//   soil::op::normal                                                             op/normal.hpp:18-39
//   soil::gradient / negslope / laplacian / slope / gaussian_blur     grad.hpp, graph.hpp:62, filter.hpp
//   soil::solve_uniform                                                          path.hpp:41-48
//   soil::albedo_layer / albedo_stratum / albedo_discharge                       erosion.hpp:110-142
//   soil::param_t                                                                erosion.hpp:17-58
//   soil::edge_t                                                                 graph.hpp:11-16
// plus soil::erode_step, the fused step the reference composes in Python (erosion_gpu.py:103-106).
//
// The reference passes silt::tensor_t<T> handles (ref-counted, by value).  silt is not part of the
// tree, so this header carries the small part of it the path needs: soil::tensor_t<T>, a
// ref-counted device buffer with a shape, which can also wrap memory the caller owns.
// "Allocating" ops return a new GPU tensor, "in-place" ops mutate their arguments - the two
// conventions of the reference.
//
// Errors: a negative status of the C ABI (bad argument, bad edge, workspace, size limit) throws
// std::invalid_argument, a positive one (cudaError_t) throws std::runtime_error - what nanobind
// turns into ValueError / RuntimeError, as the reference's own exceptions are today
// (graph.cu:76,88).  There is no CPU path: without a CUDA device every call throws.
//
// Header-only; needs the CUDA runtime API (cuda_runtime_api.h, -lcudart) and libsoilb200.so.
#pragma once

#include <array>
#include <cstddef>
#include <cstdint>
#include <initializer_list>
#include <memory>
#include <span>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include <cuda_runtime_api.h>

#include "soilb200.h"

namespace soil {

// ---- errors ---------------------------------------------------------------------------------

inline void check(int status, const char* what) {
  if (status == SOILB200_OK) return;
  if (status < 0) {
    const char* why = status == SOILB200_E_ARG         ? "invalid argument"
                      : status == SOILB200_E_EDGE      ? "unknown edge type"
                      : status == SOILB200_E_WORKSPACE ? "workspace too small"
                      : status == SOILB200_E_LIMIT     ? "map too large"
                                                       : "error";
    throw std::invalid_argument(std::string(what) + ": " + why);
  }
  throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(static_cast<cudaError_t>(status)));
}

inline void check_cuda(cudaError_t e, const char* what) {
  if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

// ---- the slice of silt the path uses ---------------------------------------------------------

using vec3 = std::array<float, 3>;

struct rng {  // silt::rng = curandStateXORWOW; opaque on the host
  unsigned char state[48];
};

class shape {
public:
  shape() = default;
  shape(std::initializer_list<int> dims) : dims_(dims) {}
  explicit shape(std::vector<int> dims) : dims_(std::move(dims)) {}
  int operator[](std::size_t d) const { return dims_.at(d); }
  std::size_t dim() const { return dims_.size(); }
  std::size_t elem() const {
    std::size_t n = 1;
    for (int d : dims_) n *= static_cast<std::size_t>(d);
    return dims_.empty() ? 0 : n;
  }
  // flatten(x, y) = x * shape[1] + y
  std::size_t flatten(int x, int y) const { return static_cast<std::size_t>(x) * dims_.at(1) + y; }
  bool operator==(const shape&) const = default;

private:
  std::vector<int> dims_;
};

template <class T>
class tensor_t {
public:
  tensor_t() = default;
  // allocate on the current device
  explicit tensor_t(soil::shape s) : shape_(std::move(s)) {
    void* p = nullptr;
    check_cuda(cudaMalloc(&p, bytes()), "tensor_t");
    data_ = std::shared_ptr<T>(static_cast<T*>(p), [](T* q) { cudaFree(q); });
  }
  // wrap device memory owned by the caller (no copy, never freed here)
  tensor_t(T* device_ptr, soil::shape s) : shape_(std::move(s)), data_(device_ptr, [](T*) {}) {}

  static tensor_t from_host(std::span<const T> host, soil::shape s) {
    tensor_t t(std::move(s));
    if (host.size() != t.elem()) throw std::invalid_argument("from_host: size does not match shape");
    check_cuda(cudaMemcpy(t.data(), host.data(), t.bytes(), cudaMemcpyHostToDevice), "from_host");
    return t;
  }
  std::vector<T> to_host() const {
    std::vector<T> out(elem());
    check_cuda(cudaMemcpy(out.data(), data(), bytes(), cudaMemcpyDeviceToHost), "to_host");
    return out;
  }
  void zero(cudaStream_t stream = nullptr) { check_cuda(cudaMemsetAsync(data(), 0, bytes(), stream), "zero"); }

  T* data() const { return data_.get(); }
  const soil::shape& shape() const { return shape_; }
  std::size_t elem() const { return shape_.elem(); }
  std::size_t bytes() const { return elem() * sizeof(T); }
  explicit operator bool() const { return static_cast<bool>(data_); }

private:
  soil::shape shape_;
  std::shared_ptr<T> data_;
};

// ---- types of the path -----------------------------------------------------------------------

enum edge_t : int { D4 = SOILB200_D4, D8 = SOILB200_D8 };  // graph.hpp:11-16

struct param_t : soilb200_param_t {  // erosion.hpp:17-58, defaults :20-56
  param_t() { soilb200_param_default(this); }
};

namespace detail {

inline void need_2d(const shape& s, const char* what) {
  if (s.dim() < 2) throw std::invalid_argument(std::string(what) + ": expected a 2-D map");
}
template <class A, class B>
void same_cells(const tensor_t<A>& a, const tensor_t<B>& b, const char* what) {
  need_2d(a.shape(), what);
  need_2d(b.shape(), what);
  if (a.shape()[0] != b.shape()[0] || a.shape()[1] != b.shape()[1])
    throw std::invalid_argument(std::string(what) + ": shape mismatch");
}
// per-device scratch that grows on demand and is reused by later calls
inline void* scratch(std::size_t bytes, int slot) {
  struct buf {
    void* p = nullptr;
    std::size_t n = 0;
  };
  static thread_local buf pool[64][3];
  int dev = 0;
  check_cuda(cudaGetDevice(&dev), "scratch");
  buf& b = pool[dev & 63][slot];
  if (b.n < bytes) {
    if (b.p) check_cuda(cudaFree(b.p), "scratch");
    b = {};
    check_cuda(cudaMalloc(&b.p, bytes), "scratch");
    b.n = bytes;
  }
  return b.p;
}

}  // namespace detail

// ---- flow graph (graph.hpp:49-60) --------------------------------------------------------------

inline tensor_t<int> direction(const tensor_t<float> height, const edge_t edge, cudaStream_t stream = nullptr) {
  detail::need_2d(height.shape(), "direction");
  tensor_t<int> out(shape{height.shape()[0], height.shape()[1]});
  check(soilb200_direction_f32(height.data(), out.data(), height.shape()[0], height.shape()[1], edge, stream),
        "direction");
  return out;
}

inline tensor_t<int> steepest(const tensor_t<float> height, const edge_t edge, cudaStream_t stream = nullptr) {
  detail::need_2d(height.shape(), "steepest");
  tensor_t<int> out(shape{height.shape()[0], height.shape()[1]});
  check(soilb200_steepest_f32(height.data(), out.data(), height.shape()[0], height.shape()[1], edge, stream),
        "steepest");
  return out;
}

// The per-cell XORWOW streams live in a cached workspace: call k with (seed, offset = k) continues the streams of
// call k-1 instead of re-running curand_init for every cell, with bit-identical draws.
inline tensor_t<int> random_weighted(const tensor_t<float> height, const edge_t edge, const std::size_t seed,
                                     const std::size_t offset, const float T, cudaStream_t stream = nullptr) {
  detail::need_2d(height.shape(), "random_weighted");
  const int nx = height.shape()[0], ny = height.shape()[1];
  tensor_t<int> out(shape{nx, ny});
  const std::size_t ws_bytes = soilb200_random_weighted_ws(nx, ny);
  void* ws = detail::scratch(ws_bytes, 0);
  check(soilb200_random_weighted_f32(height.data(), out.data(), nx, ny, edge, seed, offset, T, ws, ws_bytes, stream),
        "random_weighted");
  return out;
}

inline tensor_t<float> accumulate(const tensor_t<int> graph, const tensor_t<float> source, const edge_t edge,
                                  cudaStream_t stream = nullptr) {
  detail::same_cells(graph, source, "accumulate");
  const int nx = graph.shape()[0], ny = graph.shape()[1];
  tensor_t<float> out(shape{nx, ny});
  const std::size_t ws_bytes = soilb200_accumulate_ws(nx, ny, edge);
  check(soilb200_accumulate_f32(graph.data(), source.data(), nullptr, out.data(), nx, ny, edge,
                                detail::scratch(ws_bytes, 1), ws_bytes, stream),
        "accumulate");
  return out;
}

inline tensor_t<float> accumulate_decay(const tensor_t<int> graph, const tensor_t<float> source,
                                        const tensor_t<float> decay, const edge_t edge,
                                        cudaStream_t stream = nullptr) {
  detail::same_cells(graph, source, "accumulate_decay");
  detail::same_cells(graph, decay, "accumulate_decay");
  const int nx = graph.shape()[0], ny = graph.shape()[1];
  tensor_t<float> out(shape{nx, ny});
  const std::size_t ws_bytes = soilb200_accumulate_decay_ws(nx, ny, edge);
  check(soilb200_accumulate_f32(graph.data(), source.data(), decay.data(), out.data(), nx, ny, edge,
                                detail::scratch(ws_bytes, 1), ws_bytes, stream),
        "accumulate_decay");
  return out;
}

// The K-sample loop of example/dem_multiflow.py:43-49 on the device: float64 sum over k in
// [k_first, k_first + k_count) of accumulate(random_weighted(height, edge, seed, k, T), source) / K.
inline tensor_t<double> multiflow(const tensor_t<float> height, const tensor_t<float> source, const edge_t edge,
                                  const std::size_t K, const float T, const std::size_t seed = 0,
                                  const std::size_t k_first = 0, std::size_t k_count = SIZE_MAX,
                                  cudaStream_t stream = nullptr) {
  detail::same_cells(height, source, "multiflow");
  const int nx = height.shape()[0], ny = height.shape()[1];
  if (k_count == SIZE_MAX) k_count = K > k_first ? K - k_first : 0;
  tensor_t<double> out(shape{nx, ny});
  const std::size_t ws_bytes = soilb200_multiflow_ws(nx, ny, edge == D4 ? D4 : D8);
  check(soilb200_multiflow_f64(height.data(), source.data(), out.data(), nx, ny, edge, seed, k_first, k_count, K, T, 0,
                               detail::scratch(ws_bytes, 2), ws_bytes, stream),
        "multiflow");
  return out;
}

// ---- gradient family, slope, blur (grad.hpp, graph.hpp:62, filter.hpp) ---------------------------

using vec2 = std::array<float, 2>;

inline tensor_t<float> gradient(const tensor_t<float>& tensor, const vec2 scale, cudaStream_t stream = nullptr) {
  detail::need_2d(tensor.shape(), "gradient");
  tensor_t<float> out(shape{tensor.shape()[0], tensor.shape()[1], 2});
  check(soilb200_gradient_f32(tensor.data(), out.data(), tensor.shape()[0], tensor.shape()[1], scale.data(), stream),
        "gradient");
  return out;
}

inline tensor_t<float> negslope(const tensor_t<float>& tensor, const vec2 scale, cudaStream_t stream = nullptr) {
  detail::need_2d(tensor.shape(), "negslope");
  tensor_t<float> out(shape{tensor.shape()[0], tensor.shape()[1]});
  check(soilb200_negslope_f32(tensor.data(), out.data(), tensor.shape()[0], tensor.shape()[1], scale.data(), stream),
        "negslope");
  return out;
}

// (nx, ny, 1) or (nx, ny, 2) tensors, component-wise
inline tensor_t<float> laplacian(const tensor_t<float>& tensor, const vec2 scale, cudaStream_t stream = nullptr) {
  if (tensor.shape().dim() != 3) throw std::invalid_argument("laplacian: expected a (nx, ny, channels) tensor");
  const int nx = tensor.shape()[0], ny = tensor.shape()[1], ch = tensor.shape()[2];
  tensor_t<float> out(shape{nx, ny, ch});
  check(soilb200_laplacian_f32(tensor.data(), out.data(), nx, ny, ch, scale.data(), stream), "laplacian");
  return out;
}

inline tensor_t<float> slope(const tensor_t<float> tensor, const tensor_t<int> flow, const vec2 scale,
                             cudaStream_t stream = nullptr) {
  detail::same_cells(tensor, flow, "slope");
  tensor_t<float> out(shape{tensor.shape()[0], tensor.shape()[1]});
  check(soilb200_slope_f32(tensor.data(), flow.data(), out.data(), tensor.shape()[0], tensor.shape()[1], scale.data(),
                           stream),
        "slope");
  return out;
}

// overwrites its argument and returns it, like the reference's (filter.cu:82,90)
inline tensor_t<float> gaussian_blur(tensor_t<float> tensorIn, const float sigma, cudaStream_t stream = nullptr) {
  if (tensorIn.shape().dim() != 3) throw std::invalid_argument("gaussian_blur: expected a (nx, ny, channels) tensor");
  const int nx = tensorIn.shape()[0], ny = tensorIn.shape()[1], ch = tensorIn.shape()[2];
  tensor_t<float> tmp(shape{nx, ny, ch});
  check(soilb200_gaussian_blur_f32(tensorIn.data(), tmp.data(), nx, ny, ch, sigma, stream), "gaussian_blur");
  check_cuda(cudaStreamSynchronize(stream), "gaussian_blur");  // tmp is released on return
  return tensorIn;
}

// ---- path.hpp:41-48 ----------------------------------------------------------------------------
// Uniform-sampling Monte-Carlo estimator; one sample per generator state, `count` divides the estimate.
// Returns the (nx, ny, K) flux tensor the reference allocates (K = source.shape()[2] = 1 or 2).
inline tensor_t<float> solve_uniform(const tensor_t<float> flow, const tensor_t<float> source,
                                     const tensor_t<float> decay, tensor_t<rng> rng_states, const vec2 scale,
                                     const std::size_t count, cudaStream_t stream = nullptr) {
  if (source.shape().dim() != 3) throw std::invalid_argument("solve_uniform: source must be (nx, ny, K)");
  detail::same_cells(flow, source, "solve_uniform");
  detail::same_cells(decay, source, "solve_uniform");
  const int nx = source.shape()[0], ny = source.shape()[1], k = source.shape()[2];
  tensor_t<float> flux(shape{nx, ny, k});
  check(soilb200_solve_uniform_f32(flow.data(), source.data(), decay.data(), rng_states.data(),
                                   static_cast<int64_t>(rng_states.elem()), flux.data(), nx, ny, k, scale.data(),
                                   static_cast<int64_t>(count), stream),
        "solve_uniform");
  return flux;
}

// ---- albedo generators (erosion.hpp:110-142) ------------------------------------------------------
inline void albedo_layer(tensor_t<float> albedo, const tensor_t<float> albedoBedrock,
                         const tensor_t<float> albedoSediment, const tensor_t<float> layers,
                         const float scaleSediment, const vec3 shiftSediment, cudaStream_t stream = nullptr) {
  detail::same_cells(albedo, layers, "albedo_layer");
  check(soilb200_albedo_layer(albedo.data(), albedoBedrock.data(), albedoSediment.data(), layers.data(),
                              layers.shape()[0], layers.shape()[1], scaleSediment, shiftSediment.data(), stream),
        "albedo_layer");
}

inline void albedo_stratum(tensor_t<float> albedoBedrock, const tensor_t<float> uplift, const tensor_t<float> layers,
                           const vec3 scale, const param_t param, const vec3 colorA, const vec3 colorB,
                           const float age, const float freq, cudaStream_t stream = nullptr) {
  detail::same_cells(uplift, layers, "albedo_stratum");
  check(soilb200_albedo_stratum(albedoBedrock.data(), uplift.data(), layers.data(), layers.shape()[0],
                                layers.shape()[1], scale.data(), &param, colorA.data(), colorB.data(), age, freq,
                                stream),
        "albedo_stratum");
}

inline void albedo_discharge(tensor_t<float> albedo, const tensor_t<float> discharge, const vec3 colorDischarge,
                             const float extinction, const float scale, cudaStream_t stream = nullptr) {
  detail::same_cells(albedo, discharge, "albedo_discharge");
  check(soilb200_albedo_discharge(albedo.data(), discharge.data(), discharge.shape()[0], discharge.shape()[1],
                                  colorDischarge.data(), extinction, scale, stream),
        "albedo_discharge");
}

// silt.seed(rng, seed, offset): curand_init(seed, i, offset, &rng[i])
inline void seed(tensor_t<rng> states, const std::size_t seed_value, const std::size_t offset,
                 cudaStream_t stream = nullptr) {
  check(soilb200_seed_xorwow(states.data(), static_cast<int64_t>(states.elem()), seed_value, 0, offset, stream), "seed");
}

// ---- erosion kernels (erosion.hpp:69-142) ------------------------------------------------------

inline void transport_fluvial(tensor_t<float> layers, tensor_t<float> rainfall, tensor_t<float> discharge,
                              tensor_t<float> discharge_track, tensor_t<float> mass, tensor_t<float> mass_track,
                              tensor_t<float> momentum, tensor_t<float> momentum_track,
                              tensor_t<float> albedo_bedrock, tensor_t<float> albedo_transport,
                              tensor_t<float> albedo_surface, tensor_t<rng> rng_states, const vec3 scale,
                              const param_t param, cudaStream_t stream = nullptr) {
  detail::same_cells(layers, discharge, "transport_fluvial");
  check(soilb200_transport_fluvial(layers.data(), rainfall.data(), discharge.data(), discharge_track.data(),
                                   mass.data(), mass_track.data(), momentum.data(), momentum_track.data(),
                                   albedo_bedrock.data(), albedo_transport.data(), albedo_surface.data(),
                                   rng_states.data(), static_cast<int64_t>(rng_states.elem()), layers.shape()[0],
                                   layers.shape()[1], scale.data(), &param, stream),
        "transport_fluvial");
}

inline void transport_debris(tensor_t<float> layers, tensor_t<float> velocity, tensor_t<float> velocity_track,
                             tensor_t<float> mass, tensor_t<float> mass_track, tensor_t<float> albedo_bedrock,
                             tensor_t<float> albedo_transport, tensor_t<float> albedo_surface, tensor_t<rng> rng_states,
                             const vec3 scale, const param_t param, cudaStream_t stream = nullptr) {
  detail::same_cells(layers, mass, "transport_debris");
  check(soilb200_transport_debris(layers.data(), velocity.data(), velocity_track.data(), mass.data(),
                                  mass_track.data(), albedo_bedrock.data(), albedo_transport.data(),
                                  albedo_surface.data(), rng_states.data(), static_cast<int64_t>(rng_states.elem()),
                                  layers.shape()[0], layers.shape()[1], scale.data(), &param, stream),
        "transport_debris");
}

inline void mass_transfer(tensor_t<float> delta, tensor_t<float> layers, const tensor_t<float> uplift,
                          const tensor_t<float> discharge, const tensor_t<float> mass,
                          const tensor_t<float> momentum, const tensor_t<float> debris,
                          const tensor_t<float> momentumDebris, tensor_t<float> albedo_bedrock,
                          tensor_t<float> albedo_transport_fluvial, tensor_t<float> albedo_transport_debris,
                          tensor_t<float> albedo_surface, const vec3 scale, const param_t param,
                          cudaStream_t stream = nullptr) {
  detail::same_cells(delta, layers, "mass_transfer");
  check(soilb200_mass_transfer(delta.data(), layers.data(), uplift.data(), discharge.data(), mass.data(),
                               momentum.data(), debris.data(), momentumDebris.data(), albedo_bedrock.data(),
                               albedo_transport_fluvial.data(), albedo_transport_debris.data(), albedo_surface.data(),
                               layers.shape()[0], layers.shape()[1], scale.data(), &param, stream),
        "mass_transfer");
}

inline void mass_creep(tensor_t<float> delta, const tensor_t<float> layers, const vec3 scale, const param_t param,
                       cudaStream_t stream = nullptr) {
  detail::same_cells(delta, layers, "mass_creep");
  check(soilb200_mass_creep(delta.data(), layers.data(), layers.shape()[0], layers.shape()[1], scale.data(), &param,
                            stream),
        "mass_creep");
}

inline void layer_merge(tensor_t<float> height, const tensor_t<float> layers, cudaStream_t stream = nullptr) {
  detail::same_cells(height, layers, "layer_merge");
  check(soilb200_layer_merge(height.data(), layers.data(), layers.shape()[0], layers.shape()[1], stream), "layer_merge");
}

// ---- the fused step ----------------------------------------------------------------------------
// One call for what erosion_gpu.py:103-106 composed from the kernels above: both particle systems, their
// normalisation, transfer, creep, applying the delta and the merged height.  model_t owns the step's scratch.
class model_t {
public:
  // layers (nx, ny, 2); albedo maps (nx, ny, 3) may be empty together (no albedo channels)
  model_t(tensor_t<float> layers, tensor_t<float> rainfall, tensor_t<float> uplift, tensor_t<float> albedo_bedrock,
          tensor_t<float> albedo_surface, std::size_t particles, const vec3 scale, std::size_t seed_value = 0)
      : layers_{std::move(layers), {}}, rainfall_(std::move(rainfall)), uplift_(std::move(uplift)),
        albedo_bedrock_(std::move(albedo_bedrock)), albedo_surface_(std::move(albedo_surface)) {
    detail::need_2d(layers_[0].shape(), "model_t");
    const int nx = layers_[0].shape()[0], ny = layers_[0].shape()[1];
    const bool albedo = static_cast<bool>(albedo_bedrock_);
    if (albedo != static_cast<bool>(albedo_surface_)) throw std::invalid_argument("model_t: give both albedo maps or none");
    layers_[1] = tensor_t<float>(shape{nx, ny, 2});
    height = tensor_t<float>(shape{nx, ny});
    discharge = tensor_t<float>(shape{nx, ny});
    mass = tensor_t<float>(shape{nx, ny});
    debris = tensor_t<float>(shape{nx, ny});
    momentum = tensor_t<float>(shape{nx, ny, 2});
    debris_momentum = tensor_t<float>(shape{nx, ny, 2});
    for (auto* t : {&discharge, &mass, &debris, &momentum, &debris_momentum}) t->zero();
    rng_fluvial_ = tensor_t<rng>(shape{static_cast<int>(particles)});
    rng_debris_ = tensor_t<rng>(shape{static_cast<int>(particles)});
    seed(rng_fluvial_, seed_value, 0);
    seed(rng_debris_, seed_value + 1, 0);
    const std::size_t fb = soilb200_erode_flux_bytes(nx, ny, albedo), tb = soilb200_erode_ws(nx, ny);
    for (int i = 0; i < 2; ++i) {
      flux_[i] = tensor_t<unsigned char>(shape{static_cast<int>(fb >> 8), 256});
      tags_[i] = tensor_t<unsigned char>(shape{static_cast<int>(tb >> 8), 256});
      flux_[i].zero();
      tags_[i].zero();
    }
    pre_ = tensor_t<unsigned char>(shape{static_cast<int>(soilb200_erode_pre_bytes(nx, ny) >> 8), 256});
    mass_sum_ = tensor_t<double>(shape{2});
    scale_ = scale;
    n_particles_ = static_cast<int64_t>(particles);
    layer_merge(height, layers_[0]);
  }

  const tensor_t<float>& layers() const { return layers_[cur_]; }
  tensor_t<float> height, discharge, mass, debris, momentum, debris_momentum;

  // one erosion step; with_creep as in mass_creep's optional call
  void step(const param_t& param, bool with_creep = true, cudaStream_t stream = nullptr) {
    soilb200_model_t m{};
    m.nx = layers_[0].shape()[0];
    m.ny = layers_[0].shape()[1];
    for (int i = 0; i < 3; ++i) m.scale[i] = scale_[i];
    m.with_creep = with_creep ? 1 : 0;
    m.layers = layers_[cur_].data();
    m.layers_next = layers_[cur_ ^ 1].data();
    m.height = height.data();
    m.rainfall = rainfall_.data();
    m.uplift = uplift_.data();
    m.discharge = discharge.data();
    m.mass = mass.data();
    m.momentum = momentum.data();
    m.debris = debris.data();
    m.debris_momentum = debris_momentum.data();
    m.albedo_bedrock = albedo_bedrock_.data();
    m.albedo_surface = albedo_surface_.data();
    m.rng_fluvial = rng_fluvial_.data();
    m.rng_debris = rng_debris_.data();
    m.n_particles = n_particles_;
    m.flux = flux_[cur_].data();
    m.touched = tags_[cur_].data();
    m.flux_prev = flux_[cur_ ^ 1].data();
    m.touched_prev = tags_[cur_ ^ 1].data();
    m.mass_sum = mass_sum_.data();
    m.pre = pre_.data();
    check(soilb200_erode_step(&m, &param, stream), "erode_step");
    cur_ ^= 1;
  }

  // {sum(bedrock), sum(sediment)} after the last step, reduced in double on the device
  std::array<double, 2> mass_sums() const {
    const auto v = mass_sum_.to_host();
    return {v[0], v[1]};
  }

private:
  tensor_t<float> layers_[2];
  tensor_t<float> rainfall_, uplift_, albedo_bedrock_, albedo_surface_;
  tensor_t<rng> rng_fluvial_, rng_debris_;
  tensor_t<unsigned char> flux_[2], tags_[2], pre_;
  tensor_t<double> mass_sum_;
  vec3 scale_{};
  int64_t n_particles_ = 0;
  int cur_ = 0;
};

// ---- op/normal.hpp:18-39, op/noise.hpp:14-56 ---------------------------------------------------

namespace op {

inline tensor_t<float> normal(const tensor_t<float> height, const vec3 scale, cudaStream_t stream = nullptr) {
  detail::need_2d(height.shape(), "normal");
  tensor_t<float> out(shape{height.shape()[0], height.shape()[1], 3});
  check(soilb200_normal_f32(height.data(), out.data(), height.shape()[0], height.shape()[1], scale.data(), stream),
        "normal");
  return out;
}

}  // namespace op

struct noise_param_t {  // op/noise.hpp:14-22
  float frequency = 1.0f;
  int octaves = 8;
  float gain = 0.6f;
  float lacunarity = 2.0f;
  float seed = 0.0f;
  std::array<float, 2> ext = {512.0f, 512.0f};
};

// soil::noise (op/noise.hpp:24-56), sampled on the device
inline tensor_t<float> noise(const shape s, const noise_param_t p, cudaStream_t stream = nullptr) {
  detail::need_2d(s, "noise");
  tensor_t<float> out(shape{s[0], s[1]});
  check(soilb200_noise_f32(out.data(), s[0], s[1], p.frequency, p.octaves, p.gain, p.lacunarity, p.seed, p.ext[0],
                           p.ext[1], 1, stream),
        "noise");
  return out;
}

}  // namespace soil
