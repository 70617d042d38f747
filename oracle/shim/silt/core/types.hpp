// oracle/shim/silt - TEST INFRASTRUCTURE ONLY.
// Stand-in for the missing private submodule `silt` (ext/silt is empty,
// .SUBMODULES.json:30-36) with the conventions inferred in SURVEY.md App. A.
// It exists so that the reference's graph.cu and erosion.cu compile UNMODIFIED
// from /root/reference into oracle/_ref/ as a GPU-side oracle and baseline.
#pragma once
#include <cstddef>
#include <cmath>
#include <stdexcept>
#include <type_traits>
#include <string>
#include <cuda_runtime.h>
#include <math_constants.h>
#include <curand_kernel.h>
#include <glm/glm.hpp>

#ifndef GPU_ENABLE
#define GPU_ENABLE __host__ __device__
#endif

namespace silt {

using vec2 = glm::vec2;
using vec3 = glm::vec3;
using ivec2 = glm::ivec2;
template <int N> using fvec = glm::vec<N, float>;
template <int N> using vec = glm::vec<N, float>;  // filter.cu:80,85
using rng = curandState;  // App. A-4

enum host_t { CPU, GPU };

}  // namespace silt
