#pragma once
#include <silt/core/types.hpp>
#include <silt/core/shape.hpp>
#include <map>
#include <vector>
namespace silt {

// tiny caching device allocator so that the reference's allocate-per-call
// convention (graph.cu:80,182,540-550) does not pay cudaMalloc on every call
// when it is timed as a baseline.
struct pool {
  static std::map<size_t, std::vector<void*>>& free_list() { static std::map<size_t, std::vector<void*>> f; return f; }
  static void* get(size_t bytes) {
    auto& v = free_list()[bytes];
    if (!v.empty()) { void* p = v.back(); v.pop_back(); return p; }
    void* p = nullptr;
    if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) throw std::runtime_error("shim: cudaMalloc failed");
    return p;
  }
  // every tensor the reference allocates inside one call is tracked here and
  // handed back to the free list by recycle() once the caller has copied the
  // result out (the real silt ref-counts; the shim's handles are non-owning).
  static std::vector<std::pair<void*, size_t>>& live() { static std::vector<std::pair<void*, size_t>> l; return l; }
  static void* get_tracked(size_t bytes) { void* p = get(bytes); live().push_back({p, bytes}); return p; }
  static void recycle() { for (auto& pr : live()) free_list()[pr.second].push_back(pr.first); live().clear(); }
  static void drain() { for (auto& kv : free_list()) for (void* p : kv.second) cudaFree(p); free_list().clear(); }
};

template <typename V> struct view_t {
  V* ptr;
  GPU_ENABLE view_t() : ptr(nullptr) {}
  GPU_ENABLE view_t(V* p) : ptr(p) {}
  template <typename U, typename = std::enable_if_t<std::is_same_v<const U, V>>>
  GPU_ENABLE view_t(const view_t<U>& o) : ptr(o.ptr) {}
  GPU_ENABLE V& operator[](const size_t i) const { return ptr[i]; }
};

// non-owning device handle, trivially copyable into kernels (App. A-7)
template <typename T> struct tensor_t {
  T* ptr;
  silt::shape sh;
  host_t where;
  GPU_ENABLE tensor_t() : ptr(nullptr), sh(), where(GPU) {}
  tensor_t(const silt::shape s, const host_t h) : ptr((T*)pool::get_tracked(s.elem() * sizeof(T))), sh(s), where(h) {}
  GPU_ENABLE tensor_t(T* p, const silt::shape s) : ptr(p), sh(s), where(GPU) {}
  GPU_ENABLE T& operator[](const size_t i) const { return ptr[i]; }
  GPU_ENABLE size_t elem() const { return sh.elem(); }
  GPU_ENABLE silt::shape shape() const { return sh; }
  GPU_ENABLE host_t host() const { return where; }
  GPU_ENABLE T* data() const { return ptr; }
  template <typename V> GPU_ENABLE view_t<V> view() const { return view_t<V>((V*)ptr); }
};

// type-erased handle, only as far as path.cu needs it: built from a float tensor, read back with as<float>()
struct tensor {
  tensor_t<float> f;
  tensor() = default;
  tensor(const tensor_t<float>& t) : f(t) {}
  template <typename T> tensor_t<T> as() const { return f; }
};

}  // namespace silt
