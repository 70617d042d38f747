#pragma once
#include <silt/core/types.hpp>
namespace silt {
// App. A-2: flatten(x,y) = x*shape[1] + y, dim 1 fastest
struct shape {
  size_t d[3];
  int n;
  GPU_ENABLE shape() : d{0, 0, 0}, n(0) {}
  GPU_ENABLE shape(size_t a) : d{a, 1, 1}, n(1) {}
  GPU_ENABLE shape(size_t a, size_t b) : d{a, b, 1}, n(2) {}
  GPU_ENABLE shape(size_t a, size_t b, size_t c) : d{a, b, c}, n(3) {}
  GPU_ENABLE size_t operator[](int i) const { return d[i]; }
  GPU_ENABLE size_t elem() const { return d[0] * d[1] * d[2]; }
  GPU_ENABLE int dim() const { return n; }
  GPU_ENABLE int flatten(const ivec2 p) const { return p.x * (int)d[1] + p.y; }
  GPU_ENABLE ivec2 unflatten(const size_t i) const { return ivec2((int)(i / d[1]), (int)(i % d[1])); }
  GPU_ENABLE bool oob(const ivec2 p) const { return p.x < 0 || p.y < 0 || p.x >= (int)d[0] || p.y >= (int)d[1]; }
};
}  // namespace silt
