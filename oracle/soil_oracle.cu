// oracle/soil_oracle.cu
//
// TEST INFRASTRUCTURE ONLY.  CPU restatement of the erosiv/soillib hot path
// (reference @ 9b9e0871).  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / `--impl reference` CPU fallback may load the library built
// from this file.  Nothing under soillib_b200/ may import, link or call it.
//
// Arithmetic contract: every expression is evaluated "as written" in the
// reference source, in IEEE binary32 with round-to-nearest and NO fused
// multiply-add contraction (build with -ffp-contract=off).  The reference's
// fast intrinsics (__expf, __powf) are hardware approximations that cannot be
// reproduced on a CPU; they are restated with expf/powf and the tests carry a
// tolerance wherever a value flows through one of them.
//
// PARITY STATUS: "parity unpinned" by the reference - it ships no tests, golden
// vectors or fixtures for any function on this path.  The
// oracle is pinned instead against (1) the reference's own kernels compiled
// unmodified behind oracle/shim (oracle/_ref, runs on the GPU box; see
// tests/test_ref_gpu.py), (2) cuRAND known answers, (3) FastNoiseLite values
// generated from the reference's vendored header (tests/golden/).
//
// The file is compiled by nvcc only because the host-callable cuRAND XORWOW
// generator lives in <curand_kernel.h>; no device code is generated or run.
//
// Conventions: shape.flatten(x,y) = x*ny + y, dim 1 fastest;
// (X,Y,C) tensors are interleaved AoS; float->int conversion truncates toward
// zero with CUDA semantics (NaN -> 0, saturating).

#define QUALIFIERS static inline __host__ __device__
#include <curand_kernel.h>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <limits>
#include <omp.h>

// global scope on purpose: a parameter type with internal linkage would make the
// extern "C" functions that take it internal too
struct param_t {            // reference: erosion.hpp:17-58
  uint64_t maxage;
  float lrate, timeStep;
  float exitSlope, uplift, rainfall, gravity, evapRate;
  float frictionFactor, fluvialExponent;
  float suspensionRateFluvial, depositionRateFluvial;
  float suspensionRateDebris, depositionRateDebris, landslideRateDebris;
  float critSlopeBedrock, critSlopeSediment, yieldStress;
  float viscosityWater, bedShearWater, densityWater;
  float viscosityDebris, bedShearDebris, densityDebris;
  float force[2];
};
static_assert(sizeof(param_t) == 112, "param_t layout");

namespace {



struct v2 { float x, y; };
struct v3 { float x, y, z; };

inline v2 operator+(v2 a, v2 b) { return {a.x + b.x, a.y + b.y}; }
inline v2 operator-(v2 a) { return {-a.x, -a.y}; }
inline v2 operator*(float s, v2 a) { return {s * a.x, s * a.y}; }
inline v2 operator*(v2 a, float s) { return {a.x * s, a.y * s}; }
inline v2 operator*(v2 a, v2 b) { return {a.x * b.x, a.y * b.y}; }
inline v2 operator/(v2 a, float s) { return {a.x / s, a.y / s}; }
inline v3 operator*(float s, v3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline v3 operator+(v3 a, v3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline v3 operator/(v3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }

// sqrt(x*x + y*y): erosion_map.cu:49-53, graph.cu:23-25, glm::length
inline float len2(v2 v) { return sqrtf(v.x * v.x + v.y * v.y); }

// CUDA cvt.rzi.s32.f32 semantics (x86 would give INT_MIN for NaN)
inline int f2i(float f) {
  if (f != f) return 0;
  if (f >= 2147483648.0f) return std::numeric_limits<int>::max();
  if (f <= -2147483648.0f) return std::numeric_limits<int>::min();
  return (int)f;
}

struct grid_t {
  int nx, ny;
  inline bool oob(int x, int y) const { return x < 0 || y < 0 || x >= nx || y >= ny; }
  inline int flat(int x, int y) const { return x * ny + y; }
  inline size_t elem() const { return (size_t)nx * (size_t)ny; }
};

// neighbour tables: graph.hpp:21-46
const int SHIFT[8][2] = {{-1, 0}, {0, -1}, {0, 1}, {1, 0}, {-1, -1}, {-1, 1}, {1, -1}, {1, 1}};
inline int edgeK(int edge) { return edge == 0 ? 4 : 8; }
inline float shiftlen(int k) { return sqrtf((float)(SHIFT[k][0] * SHIFT[k][0]) + (float)(SHIFT[k][1] * SHIFT[k][1])); }

int g_threads = 0;
inline int nthreads() { return g_threads > 0 ? g_threads : omp_get_max_threads(); }

// ---------------------------------------------------------------------------
// erosion_map.cu helpers
// ---------------------------------------------------------------------------

inline bool oobf(const grid_t& g, v2 p) {  // erosion_map.cu:29-40
  if (p.x < 0) return true;
  if (p.y < 0) return true;
  if (p.x >= g.nx) return true;
  if (p.y >= g.ny) return true;
  return false;
}

inline int flattenf(const grid_t& g, v2 p) {  // erosion_map.cu:42-47 (ivec2 truncation, App. A-5)
  return f2i(p.x) * g.ny + f2i(p.y);
}

inline float stepsize(v2 p, v2 d) {  // erosion_map.cu:56-78
  const float tmax = 1.41421356237309504880f;  // CUDART_SQRT_TWO_F
  const float x_neg = floorf(p.x);
  const float y_neg = floorf(p.y);
  const float x_pos = 1.0f + x_neg;
  const float y_pos = 1.0f + y_neg;
  const float tx_neg = (x_neg - p.x) / d.x;
  const float tx_pos = (x_pos - p.x) / d.x;
  const float tx = fminf(fmaxf(tx_neg, tx_pos), tmax);
  const float ty_neg = (y_neg - p.y) / d.y;
  const float ty_pos = (y_pos - p.y) / d.y;
  const float ty = fminf(fmaxf(ty_neg, ty_pos), tmax);
  return 0.5f * (tx + ty);
}

inline float heightAt(const v2* layers, long i) { return layers[i].x + layers[i].y; }  // erosion_map.cu:99-105

// erosion_map.cu:107-159 (__grad just forwards, :161-171)
inline v2 glocal(const v2* layers, const grid_t& g, const float* s, int px, int py, float exitSlope) {
  const float nan = std::numeric_limits<float>::quiet_NaN();
  const float h = heightAt(layers, (long)px * g.ny + py);
  const float hn0 = g.oob(px - 1, py) ? nan : heightAt(layers, (long)(px - 1) * g.ny + py);
  const float hp0 = g.oob(px + 1, py) ? nan : heightAt(layers, (long)(px + 1) * g.ny + py);
  const float h0n = g.oob(px, py - 1) ? nan : heightAt(layers, (long)px * g.ny + (py - 1));
  const float h0p = g.oob(px, py + 1) ? nan : heightAt(layers, (long)px * g.ny + (py + 1));

  float gxn = (h - hn0) * s[2] / s[0];
  if (gxn != gxn) gxn = exitSlope; else gxn = fmaxf(gxn, 0.0f);
  float gyn = (h - h0n) * s[2] / s[1];
  if (gyn != gyn) gyn = exitSlope; else gyn = fmaxf(gyn, 0.0f);
  float gxp = (hp0 - h) * s[2] / s[0];
  if (gxp != gxp) gxp = -exitSlope; else gxp = fminf(gxp, 0.0f);
  float gyp = (h0p - h) * s[2] / s[1];
  if (gyp != gyp) gyp = -exitSlope; else gyp = fminf(gyp, 0.0f);

  float gx = 0.0f;
  if (fabsf(gxn) > fabsf(gx)) gx = gxn;
  if (fabsf(gxp) > fabsf(gx)) gx = gxp;
  float gy = 0.0f;
  if (fabsf(gyn) > fabsf(gy)) gy = gyn;
  if (fabsf(gyp) > fabsf(gy)) gy = gyp;
  return {gx, gy};
}

inline v2 gradAt(const v2* layers, const grid_t& g, const float* s, v2 pos, float exitSlope) {
  return glocal(layers, g, s, f2i(pos.x), f2i(pos.y), exitSlope);
}

inline void atomic_addf(float* p, float v) {
#pragma omp atomic
  *p += v;
}

inline v3 mmin(float f, v3 v) { return {fminf(v.x, f), fminf(v.y, f), fminf(v.z, f)}; }  // erosion_map.cu:88-97

}  // namespace

extern "C" {

void orc_set_threads(int n) { g_threads = n; }
int orc_get_threads() { return nthreads(); }

// ---------------------------------------------------------------------------
// graph.cu:27-70 (steepest) and :201-243 (direction)
// ---------------------------------------------------------------------------

static void steepest_impl(const float* h, int* out, int nx, int ny, int edge, bool code) {
  const grid_t g{nx, ny};
  const int K = edgeK(edge);
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const float hl = h[g.flat(x, y)];
      float smax = 0.0f;
      int next = -1;
      for (int k = 0; k < K; ++k) {
        const int qx = x + SHIFT[k][0], qy = y + SHIFT[k][1];
        if (g.oob(qx, qy)) continue;
        const int nind = g.flat(qx, qy);
        const float scur = (hl - h[nind]) / shiftlen(k);
        if (scur > smax) { smax = scur; next = code ? k : nind; }
      }
      out[g.flat(x, y)] = next;
    }
}

void orc_steepest(const float* h, int* out, int nx, int ny, int edge) { steepest_impl(h, out, nx, ny, edge, false); }
void orc_direction(const float* h, int* out, int nx, int ny, int edge) { steepest_impl(h, out, nx, ny, edge, true); }

// ---------------------------------------------------------------------------
// RNG: silt::rng == curandStateXORWOW (48 B), App. A-4; graph.cu:97-101
// ---------------------------------------------------------------------------

int orc_rng_state_bytes() { return (int)sizeof(curandStateXORWOW_t); }

void orc_seed(void* states, long n, unsigned long long seed, unsigned long long offset) {
  curandStateXORWOW_t* st = (curandStateXORWOW_t*)states;
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) curand_init(seed, (unsigned long long)i, offset, &st[i]);
}

// subsequence base lets tests seed a window [first, first+n) of a larger tensor
void orc_seed_range(void* states, long first, long n, unsigned long long seed, unsigned long long offset) {
  curandStateXORWOW_t* st = (curandStateXORWOW_t*)states;
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) curand_init(seed, (unsigned long long)(first + i), offset, &st[i]);
}

void orc_uniform(void* states, long n, float* out) {
  curandStateXORWOW_t* st = (curandStateXORWOW_t*)states;
  for (long i = 0; i < n; ++i) out[i] = curand_uniform(&st[i]);
}

// ---------------------------------------------------------------------------
// graph.cu:103-195 random_weighted.  `ambiguous` (optional) is set to 1 where
// |u - CDF_k/Z| is within `margin` of a decision boundary, i.e. where the
// reference's ex2.approx based __expf may legitimately pick the other branch.
// ---------------------------------------------------------------------------

void orc_random_weighted(const float* h, int* out, int nx, int ny, int edge, unsigned long long seed,
                         unsigned long long offset, float T, unsigned char* ambiguous, float margin) {
  const grid_t g{nx, ny};
  const int K = edgeK(edge);
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const int n = g.flat(x, y);
      curandStateXORWOW_t st;
      curand_init(seed, (unsigned long long)n, offset, &st);
      const float hl = h[n];
      float CDF[8];
      float Z = 0.0f;
      for (int k = 0; k < K; ++k) {
        const int qx = x + SHIFT[k][0], qy = y + SHIFT[k][1];
        if (g.oob(qx, qy)) continue;
        const float dE = (hl - h[g.flat(qx, qy)]) / shiftlen(k);
        const float P = (dE <= 0.0f) ? 0.0f : expf(dE / T);
        CDF[k] = Z + P;
        Z += P;
      }
      int next = -1;
      const float u = curand_uniform(&st);
      bool amb = false;
      for (int k = 0; k < K; ++k) {
        const int qx = x + SHIFT[k][0], qy = y + SHIFT[k][1];
        if (g.oob(qx, qy)) continue;
        const float c = CDF[k] / Z;
        if (fabsf(u - c) <= margin) amb = true;
        if (next < 0 && u < c) next = g.flat(qx, qy);
      }
      out[n] = next;
      if (ambiguous) ambiguous[n] = amb ? 1 : 0;
    }
}

// ---------------------------------------------------------------------------
// graph.cu:321-576 accumulate / accumulate_decay.
// mode 0: the reference algorithm including its fixed round schedule.
// mode 1: exact upstream sum in topological order; at each cell donors are
//         added in slot order k=0..K-1 (the compacted order of __count).
// decay == nullptr -> scalar decay (accumulate uses 1.0f).
// ---------------------------------------------------------------------------

static void build_donors(const int* graph, const grid_t& g, int K, std::vector<int>& donor, std::vector<int>& count) {
  const size_t N = g.elem();
  donor.assign(N * K, -1);
  count.assign(N, 0);
  // __donor graph.cu:321-348: slot k of the receiver is unique per (recv, k)
  for (int x = 0; x < g.nx; ++x)
    for (int y = 0; y < g.ny; ++y) {
      const int n = g.flat(x, y);
      const int next = graph[n];
      for (int k = 0; k < K; ++k) {
        const int qx = x + SHIFT[k][0], qy = y + SHIFT[k][1];
        if (g.oob(qx, qy)) continue;
        const int nind = g.flat(qx, qy);
        if (nind == next) donor[(size_t)K * nind + k] = n;
      }
    }
  // __count graph.cu:350-380: compact to the front
  for (size_t n = 0; n < N; ++n) {
    int c = 0;
    int tmp[8];
    for (int k = 0; k < K; ++k) {
      const int d = donor[K * n + k];
      if (d >= 0) tmp[c++] = d;
    }
    count[n] = c;
    for (int k = 0; k < K; ++k) donor[K * n + k] = (k < c) ? tmp[k] : -1;
  }
}

void orc_accumulate(const int* graph, const float* src, const float* decay, float decay_scalar, float* out, int nx,
                    int ny, int edge, int mode, int* rounds_to_converge) {
  const grid_t g{nx, ny};
  const int K = edgeK(edge);
  const size_t N = g.elem();
  std::vector<int> donor, count;
  build_donors(graph, g, K, donor, count);
  // my_decay graph.cu:382-420: keyed on the COMPACTED slot index (App. B-7)
  std::vector<float> w(N * K, 0.0f);
  for (size_t n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) {
      const int d = donor[K * n + k];
      if (d < 0) break;
      const float din = decay ? decay[d] : decay_scalar;
      w[K * n + k] = (k < 4) ? din : powf(din, 1.414f);
    }

  if (mode == 1) {
    // exact: Kahn order from leaves; value[n] = src[n] + sum_k w_k * value[donor_k]
    std::vector<int> pending(count);
    std::vector<int> stack;
    stack.reserve(N);
    for (size_t n = 0; n < N; ++n) {
      out[n] = src[n];
      if (count[n] == 0) stack.push_back((int)n);
    }
    std::vector<char> done(N, 0);
    while (!stack.empty()) {
      const int n = stack.back();
      stack.pop_back();
      float v = src[n];
      for (int k = 0; k < count[n]; ++k) v += w[(size_t)K * n + k] * out[donor[(size_t)K * n + k]];
      out[n] = v;
      done[n] = 1;
      const int r = graph[n];
      if (r < 0 || (size_t)r >= N) continue;
      // only counts if n is registered as donor of r (neighbour test in __donor)
      bool reg = false;
      for (int k = 0; k < count[r]; ++k) reg |= (donor[(size_t)K * r + k] == n);
      if (reg && --pending[r] == 0) stack.push_back(r);
    }
    if (rounds_to_converge) *rounds_to_converge = 0;
    return;
  }

  // reference schedule: graph.cu:429-563
  struct acc { std::vector<int> donor, count; std::vector<float> value, decay; };
  acc A{donor, count, std::vector<float>(src, src + N), w};
  acc B{std::vector<int>(N * K, -1), std::vector<int>(N, 0), std::vector<float>(N, 0.0f), std::vector<float>(N * K, 0.0f)};
  const size_t iter = (size_t)std::ceil(std::log2f((float)N) / 2.0f);
  auto round = [&](acc& o, const acc& in) {
    long active = 0;
#pragma omp parallel for num_threads(nthreads()) schedule(static) reduction(+ : active)
    for (long n = 0; n < (long)N; ++n) {
      float value = in.value[n];
      int cnt = in.count[n];
      int donors[8];
      float decays[8];
      for (int k = 0; k < cnt; ++k) { donors[k] = in.donor[(size_t)K * n + k]; decays[k] = in.decay[(size_t)K * n + k]; }
      for (int k = 0; k < cnt; ++k) {
        const int d = donors[k];
        const float dc = decays[k];
        const int dcount = in.count[d];
        if (dcount == 0) {
          value += dc * in.value[d];
          donors[k] = donors[cnt - 1];
          decays[k] = decays[cnt - 1];
          donors[cnt - 1] = -1;
          decays[cnt - 1] = 0.0f;
          cnt -= 1;
          k -= 1;
        } else if (dcount == 1) {
          value += dc * in.value[d];
          donors[k] = in.donor[(size_t)K * d];
          decays[k] = dc * in.decay[(size_t)K * d];
        }
      }
      o.value[n] = value;
      o.count[n] = cnt;
      for (int k = 0; k < cnt; ++k) { o.donor[(size_t)K * n + k] = donors[k]; o.decay[(size_t)K * n + k] = decays[k]; }
      active += cnt;
    }
    return active;
  };
  int conv = -1, r = 0;
  for (size_t i = 0; i <= iter; ++i) {
    long a = round(B, A); ++r;
    if (a == 0 && conv < 0) conv = r;
    a = round(A, B); ++r;
    if (a == 0 && conv < 0) conv = r;
  }
  if (rounds_to_converge) *rounds_to_converge = conv;
  std::memcpy(out, A.value.data(), N * sizeof(float));
}

// ---------------------------------------------------------------------------
// erosion.cu:29-239 transport_fluvial (particles + normalise)
// fields follow the header order erosion.hpp:69-84
// ---------------------------------------------------------------------------

void orc_transport_fluvial(const float* layers_, const float* rainfall, float* waterHeight, float* waterFlux,
                           float* mass, float* massFlux, float* momentum_, float* velocityFlux_,
                           const float* /*albedo_bedrock*/, float* albedoFlux_, const float* albedoSource_,
                           void* rng_, long P, int nx, int ny, const float* scale, const param_t* prm,
                           int do_particles, int do_normalize) {
  const grid_t g{nx, ny};
  const v2* layers = (const v2*)layers_;
  v2* velocity = (v2*)momentum_;
  v2* velocityFlux = (v2*)velocityFlux_;
  v3* albedoFlux = (v3*)albedoFlux_;
  const v3* albedoSource = (const v3*)albedoSource_;
  curandStateXORWOW_t* rng = (curandStateXORWOW_t*)rng_;
  const param_t param = *prm;
  const size_t elem = g.elem();

  if (do_particles) {
#pragma omp parallel for num_threads(nthreads()) schedule(dynamic, 16)
    for (long n = 0; n < P; ++n) {
      const float A = scale[0] * scale[1];
      const v2 L{scale[0], scale[1]};
      const float Pp = 1.0f / float(A * (float)elem);
      const float Q = 1.0f / (Pp * (float)(size_t)P);
      const float eps = 1E-12f;
      v2 pos;
      pos.x = 0.5f + curand_uniform(&rng[n]) * float(nx - 1);
      pos.y = 0.5f + curand_uniform(&rng[n]) * float(ny - 1);
      int ind = flattenf(g, pos);

      const float rho_w = param.densityWater;
      const float tau = param.bedShearWater;
      const float nu = param.viscosityWater;
      const float gr = param.gravity;
      const float ks = param.suspensionRateFluvial / 64.0f;
      const float kd = param.depositionRateFluvial * 1.33f;
      const float fD = param.frictionFactor / 8.0f;
      const float alpha = param.fluvialExponent;
      const float R = param.rainfall;
      const v2 force{param.force[0], param.force[1]};

      const v2 vel = velocity[ind];
      v2 grad = gradAt(layers, g, scale, pos, param.exitSlope);
      v2 speed = -(gr * grad) + nu * vel + force;
      speed = speed / sqrtf(len2(L * speed));
      if (len2(speed) < eps) continue;

      const float v = len2(vel);
      const float shear = 0.125f * fD * rho_w * v * v;
      const float power = powf(shear * len2(grad), alpha);  // __powf on the GPU

      const float source_m = Q * ks * power;
      const float source_w = Q * R * rainfall[ind];
      const v2 source_v = Q * (-(gr * grad) + nu * vel);
      const v3 source_a = source_m * albedoSource[ind];

      float att_w = 1.0f, att_m = 1.0f, att_v = 1.0f;
      int iter = 0;
      while (!oobf(g, pos) && (uint64_t)(++iter) < param.maxage) {
        const int nind = flattenf(g, pos);
        if (nind != ind) {
          ind = nind;
          atomic_addf(&waterFlux[ind], att_w * source_w);
          atomic_addf(&massFlux[ind], att_m * source_m);
          atomic_addf(&velocityFlux[ind].x, att_v * source_v.x);
          atomic_addf(&velocityFlux[ind].y, att_v * source_v.y);
          atomic_addf(&albedoFlux[ind].x, att_m * source_a.x);
          atomic_addf(&albedoFlux[ind].y, att_m * source_a.y);
          atomic_addf(&albedoFlux[ind].z, att_m * source_a.z);
        }
        const float v_norm = len2(speed);
        const v2 v_unit = speed / v_norm;
        const float v_step = stepsize(pos, v_unit);
        const float dL = v_step * len2(L);
        const float ds = dL / v_norm;
        if (v_norm < eps) break;

        grad = gradAt(layers, g, scale, pos, param.exitSlope);
        const v2 accel = -(gr * grad) + nu * velocity[ind] + force;
        speed = (1.0f / (1.0f + dL * (tau + nu))) * speed + (dL / (1.0f + dL * (tau + nu))) * accel;

        const float decay_m = kd;
        const float decay_w = param.evapRate;
        const float decay_v = 0.125f * fD / (eps + waterHeight[ind]);
        att_m = att_m * expf(-ds * decay_m);  // __expf on the GPU
        att_w = att_w * expf(-ds * decay_w);
        att_v = att_v * expf(-dL * decay_v);
        pos = pos + v_step * v_unit;
      }
    }
  }

  if (do_normalize) {  // erosion.cu:143-187
#pragma omp parallel for num_threads(nthreads()) schedule(static)
    for (long n = 0; n < (long)elem; ++n) {
      const float A = scale[0] * scale[1];
      const v2 L{scale[0], scale[1]};
      const v2 vv{1.0f, 0.0f};
      const float norm = fabsf(vv.x * L.y) + fabsf(vv.y * L.x);
      const int px = (int)(n / ny), py = (int)(n % ny);
      const v2 grad = glocal(layers, g, scale, px, py, param.exitSlope);
      const float m = massFlux[n];
      const v3 a = albedoFlux[n];
      const float source_w = param.rainfall * rainfall[n];
      const v2 source_v = -param.gravity * grad + v2{param.force[0], param.force[1]};
      const float source_m = 0.0f;
      waterHeight[n] = (A * source_w + waterFlux[n]) / norm;
      mass[n] = (A * source_m + m) / norm;
      velocity[n] = (A * source_v + velocityFlux[n]) / norm;
      if (m > 0.0f && len2(v2{a.x, a.y}) > 0.0f) albedoFlux[n] = a / m;  // App. B-5: vec3 -> vec2 truncation
      else albedoFlux[n] = albedoSource[n];
    }
  }
}

// ---------------------------------------------------------------------------
// erosion.cu:245-436 transport_debris (erosion.hpp:86-98 order)
// ---------------------------------------------------------------------------

void orc_transport_debris(const float* layers_, float* velocity_, float* velocityFlux_, float* mass, float* massFlux,
                          const float* /*albedo_bedrock*/, float* albedoFlux_, const float* albedoSource_, void* rng_,
                          long P, int nx, int ny, const float* scale, const param_t* prm, int do_particles,
                          int do_normalize) {
  const grid_t g{nx, ny};
  const v2* layers = (const v2*)layers_;
  v2* velocity = (v2*)velocity_;
  v2* velocityFlux = (v2*)velocityFlux_;
  v3* albedoFlux = (v3*)albedoFlux_;
  const v3* albedoSource = (const v3*)albedoSource_;
  curandStateXORWOW_t* rng = (curandStateXORWOW_t*)rng_;
  const param_t param = *prm;
  const size_t elem = g.elem();

  if (do_particles) {
#pragma omp parallel for num_threads(nthreads()) schedule(dynamic, 16)
    for (long n = 0; n < P; ++n) {
      const float A = scale[0] * scale[1];
      const v2 L{scale[0], scale[1]};
      const float Pp = 1.0f / float(A * (float)elem);
      const float Q = 1.0f / (Pp * (float)(size_t)P);
      const float eps = 1E-12f;
      v2 pos;
      pos.x = 0.5f + curand_uniform(&rng[n]) * float(nx - 1);
      pos.y = 0.5f + curand_uniform(&rng[n]) * float(ny - 1);
      int ind = flattenf(g, pos);

      const float theta = param.critSlopeBedrock;
      const float nu = param.viscosityDebris;
      const float tau = param.bedShearDebris;
      const float gr = param.gravity;
      const float kl = param.landslideRateDebris;
      const float kdd = param.depositionRateDebris;
      const float kds = param.suspensionRateDebris;
      const float tau_y = param.yieldStress;

      const v2 vel = velocity[ind];
      v2 grad = gradAt(layers, g, scale, pos, param.exitSlope);
      v2 speed = -(gr * grad) + nu * vel;
      speed = speed / sqrtf(len2(L * speed));
      if (len2(speed) < eps) continue;

      const float excessSlope0 = (len2(grad) - theta);
      const float suspend = fmaxf(0.0f, kl * excessSlope0);
      const float source_d = Q * suspend;
      const v2 source_v = Q * (-gr * grad + nu * vel);
      const v3 source_a = source_d * albedoSource[ind];

      float att_d = 1.0f, att_v = 1.0f;
      int iter = 0;
      while (!oobf(g, pos) && (uint64_t)(++iter) < param.maxage) {
        const int nind = flattenf(g, pos);
        if (nind != ind) {
          ind = nind;
          atomic_addf(&massFlux[ind], att_d * source_d);
          atomic_addf(&velocityFlux[ind].x, att_v * source_v.x);
          atomic_addf(&velocityFlux[ind].y, att_v * source_v.y);
          atomic_addf(&albedoFlux[ind].x, att_d * source_a.x);
          atomic_addf(&albedoFlux[ind].y, att_d * source_a.y);
          atomic_addf(&albedoFlux[ind].z, att_d * source_a.z);
        }
        const float v_norm = len2(speed);
        const v2 v_unit = speed / v_norm;
        const float v_step = stepsize(pos, v_unit);
        const float dL = v_step * len2(L);
        const float ds = dL / v_norm;
        if (v_norm < eps) break;

        grad = gradAt(layers, g, scale, pos, param.exitSlope);
        const float debrisHeight = eps + att_d * source_d;
        const v2 accel = -(gr * grad) + nu * velocity[ind];
        const float decay = nu + tau / debrisHeight;
        const float w = 1.0f / (1.0f + dL * decay);
        speed = w * speed + w * dL * accel;

        const float excessSlope = (len2(grad) - theta);
        const float excessStress = gr * (excessSlope - tau_y / debrisHeight);
        const float shearRate = (excessStress < 0.0f) ? kdd : kds;
        const float decay_d = ds * shearRate * excessStress / v_norm;
        const float decay_v = (nu + tau / debrisHeight);
        att_d = att_d * expf(decay_d);
        att_v = att_v * expf(-dL * decay_v);
        pos = pos + v_step * v_unit;
      }
    }
  }

  if (do_normalize) {  // erosion.cu:353-393
#pragma omp parallel for num_threads(nthreads()) schedule(static)
    for (long n = 0; n < (long)elem; ++n) {
      const float A = scale[0] * scale[1];
      const v2 L{scale[0], scale[1]};
      const v2 vv{1.0f, 0.0f};
      const float norm = fabsf(vv.x * L.y) + fabsf(vv.y * L.x);
      const int px = (int)(n / ny), py = (int)(n % ny);
      const v2 grad = glocal(layers, g, scale, px, py, param.exitSlope);
      const float m = massFlux[n];
      const v3 a = albedoFlux[n];
      const v2 source_v = -param.gravity * grad;
      const float source_d = 0.0f;
      mass[n] = (A * source_d + m) / norm;
      velocity[n] = (A * source_v + velocityFlux[n]) / norm;
      if (m > 0.0f && len2(v2{a.x, a.y}) > 0.0f) albedoFlux[n] = a / m;
      else albedoFlux[n] = albedoSource[n];
    }
  }
}

// ---------------------------------------------------------------------------
// erosion.cu:453-611 mass_transfer (erosion.hpp:104-119 order).  albedo
// pointers may all be null (our optional-albedo extension; the reference
// always has them).
// ---------------------------------------------------------------------------

void orc_mass_transfer(float* delta_, const float* layers_, const float* upliftBase, const float* /*waterHeight*/,
                       const float* mass, const float* velocityFluvial_, const float* debris,
                       const float* /*momentumDebris*/, const float* albedo_bedrock_, const float* albedoFluvial_,
                       const float* albedoDebris_, float* albedo_surface_, int nx, int ny, const float* scale,
                       const param_t* prm) {
  const grid_t g{nx, ny};
  v2* deltas = (v2*)delta_;
  const v2* layers = (const v2*)layers_;
  const v2* velocityFluvial = (const v2*)velocityFluvial_;
  const v3* albedo_bedrock = (const v3*)albedo_bedrock_;
  const v3* albedoFluxFluvial = (const v3*)albedoFluvial_;
  const v3* albedoFluxDebris = (const v3*)albedoDebris_;
  v3* albedo_surface = (v3*)albedo_surface_;
  const param_t param = *prm;
  const size_t elem = g.elem();
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long n = 0; n < (long)elem; ++n) {
    const float dt = param.timeStep;
    const float ku = param.uplift;
    const float kfs = param.suspensionRateFluvial / 64.0f;
    const float kfd = param.depositionRateFluvial * 1.33f;
    const float fD = param.frictionFactor / 8.0f;
    const float alpha = param.fluvialExponent;
    const float rho = param.densityWater;
    const float gr = param.gravity;
    const float tau_y = param.yieldStress;
    const float kds = param.suspensionRateDebris;
    const float kdd = param.depositionRateDebris;
    const float kL = param.landslideRateDebris;
    const float eps = 1E-12f;

    const int px = (int)(n / ny), py = (int)(n % ny);
    const v2 grad = glocal(layers, g, scale, px, py, param.exitSlope);
    const float L = len2(v2{scale[0], scale[1]});
    const float slope = len2(grad);

    const v2 speed = velocityFluvial[n];
    const float v = len2(speed);
    const float shear = 0.125f * fD * rho * v * v;
    const float power = powf(shear * slope, alpha);  // __powf on the GPU
    const float suspend = kfs * power;

    const float massHeight = mass[n];
    const float deposit = kfd * massHeight;
    const float uplift = ku * upliftBase[n];

    const float debrisHeight = debris[n];
    const float excessSlope = (slope - param.critSlopeBedrock);
    const float shearLandslide = fmaxf(0.0f, kL * excessSlope);
    const float shearYield = gr * (debrisHeight * excessSlope - tau_y);
    const float suspendDebris = shearLandslide + kds * fmaxf(0.0f, shearYield);
    const float depositDebris = fminf(debrisHeight, fmaxf(0.0f, -kdd * shearYield));

    float transfer = dt * (deposit - suspend + depositDebris - suspendDebris);
    transfer = fmaxf(transfer, -0.25f * L * slope);
    transfer = fminf(transfer, 0.25f * L * 0.3f);

    const v2 layer = layers[n];
    v2 delta = deltas[n];
    delta.x += dt * uplift / scale[2];
    delta.y += fmaxf(0.0f, transfer / scale[2]);
    if (transfer < 0.0f) {
      const float limited = fmaxf(-layer.y * scale[2], transfer);
      delta.y += limited / scale[2];
      transfer -= limited;
      delta.x += transfer / scale[2];
    }
    deltas[n] = delta;

    if (!albedo_surface) continue;
    const v3 albedoFluvial = albedoFluxFluvial[n];
    const v3 albedoDebris = albedoFluxDebris[n];
    const float totalHeight = massHeight + debrisHeight;
    const float mixDepth = 1.0f;
    if (layer.y == 0.0f) {
      albedo_surface[n] = albedo_bedrock[n];
    } else if (totalHeight > 0.0f && transfer > eps) {
      const float wMass = fminf(massHeight / totalHeight, 1.0f);
      const v3 colorTransport = mmin(1.0f, wMass * albedoFluvial + (1.0f - wMass) * albedoDebris);
      const v3 colorSurface = mmin(1.0f, albedo_surface[n]);
      const float wSurf = fminf(mixDepth, layer.y * scale[2]);
      const float wTrsp = fmaxf(eps, transfer);
      const float w = fminf(wTrsp / (wTrsp + wSurf), 1.0f);
      albedo_surface[n] = w * colorTransport + (1.0f - w) * colorSurface;
    }
  }
}

// ---------------------------------------------------------------------------
// erosion.cu:633-727 mass_creep
// ---------------------------------------------------------------------------

void orc_mass_creep(float* delta_, const float* layers_, int nx, int ny, const float* scale, const param_t* prm) {
  const grid_t g{nx, ny};
  v2* delta = (v2*)delta_;
  const v2* layers = (const v2*)layers_;
  const float critSlope = prm->critSlopeSediment;
  const float sz = scale[2];
  auto tr = [critSlope, sz](const v2& lb, const v2& lt, const float dx) {
    const float hb = (lb.x + lb.y) * sz;
    const float ht = (lt.x + lt.y) * sz;
    const float tmax = 0.5f * ((ht - hb) - critSlope * dx);
    return fmaxf(0.0f, fminf(lt.y * sz, tmax));
  };
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const int n = g.flat(x, y);
      const v2 l00 = layers[n];
      const v2 ln0 = g.oob(x - 1, y) ? l00 : layers[g.flat(x - 1, y)];
      const v2 lp0 = g.oob(x + 1, y) ? l00 : layers[g.flat(x + 1, y)];
      const v2 l0n = g.oob(x, y - 1) ? l00 : layers[g.flat(x, y - 1)];
      const v2 l0p = g.oob(x, y + 1) ? l00 : layers[g.flat(x, y + 1)];
      const float h00 = (l00.x + l00.y) * sz;
      const float hn0 = (ln0.x + ln0.y) * sz;
      const float hp0 = (lp0.x + lp0.y) * sz;
      const float h0n = (l0n.x + l0n.y) * sz;
      const float h0p = (l0p.x + l0p.y) * sz;
      float t = 0.0f;
      if (hp0 > h00) t += tr(l00, lp0, scale[0]); else t -= tr(lp0, l00, scale[0]);
      if (hn0 > h00) t += tr(l00, ln0, scale[0]); else t -= tr(ln0, l00, scale[0]);
      if (h0p > h00) t += tr(l00, l0p, scale[1]); else t -= tr(l0p, l00, scale[1]);
      if (h0n > h00) t += tr(l00, l0n, scale[1]); else t -= tr(l0n, l00, scale[1]);
      delta[n].y += 0.25f * t / sz;
    }
}

// erosion.cu:733-757
void orc_layer_merge(float* height, const float* layers_, long n) {
  const v2* layers = (const v2*)layers_;
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) height[i] = layers[i].x + layers[i].y;
}

// silt.add(layers, delta): the INFERRED apply step
void orc_add(float* a, const float* b, long n) {
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) a[i] = a[i] + b[i];
}

// ---------------------------------------------------------------------------
// op/normal.hpp:18-39.  The gradient stencil (silt lerp5_t::grad) is not in
// the tree; SURVEY.md App. A-8 specifies it: per axis 4th-order central
// difference when f(-2..+2) all exist (in bounds, not NaN), else 2nd-order
// central, else one-sided, else 0 (the constant divisions written as multiplications by 1/12 and 1/2);
// gradient = d * (scale.z / scale.xy), the ratio rounded once.
// "spec-defined, parity unpinned".
// ---------------------------------------------------------------------------

static inline float axis_grad(float fm2, float fm1, float f0, float fp1, float fp2) {
  const bool m2 = fm2 == fm2, m1 = fm1 == fm1, p1 = fp1 == fp1, p2 = fp2 == fp2;
  if (f0 != f0) return 0.0f;
  if (m2 && m1 && p1 && p2) return ((fm2 - 8.0f * fm1) + (8.0f * fp1 - fp2)) * (1.0f / 12.0f);
  if (m1 && p1) return (fp1 - fm1) * 0.5f;
  if (p1) return fp1 - f0;
  if (m1) return f0 - fm1;
  return 0.0f;
}

void orc_normal(const float* h, float* out, int nx, int ny, const float* scale) {
  const grid_t g{nx, ny};
  const float nan = std::numeric_limits<float>::quiet_NaN();
  auto at = [&](int x, int y) { return g.oob(x, y) ? nan : h[g.flat(x, y)]; };
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const float f0 = at(x, y);
      const float dx = axis_grad(at(x - 2, y), at(x - 1, y), f0, at(x + 1, y), at(x + 2, y));
      const float dy = axis_grad(at(x, y - 2), at(x, y - 1), f0, at(x, y + 1), at(x, y + 2));
      const float rx = scale[2] / scale[0], ry = scale[2] / scale[1];
      const float nxv = -(dx * rx), nyv = -(dy * ry), nzv = 1.0f;
      const float inv = 1.0f / sqrtf(nxv * nxv + nyv * nyv + nzv * nzv);  // glm::normalize = v * inversesqrt(dot)
      float* o = out + 3 * (size_t)g.flat(x, y);
      o[0] = nxv * inv;
      o[1] = nyv * inv;
      o[2] = nzv * inv;
    }
}

// ---------------------------------------------------------------------------
// op/noise.hpp:14-56 over FastNoiseLite 1.0.1 (vendored at
// source/soillib/external/FastNoiseLite.h): OpenSimplex2 3-D, FBm.
// Restated from the published algorithm: GetNoise :322-336, coordinate
// rotation :687-727, FBm :866-886, fractal bounding :473-484, hash/gradient
// :500-553, SingleOpenSimplex2 :1054-1148.  Pinned by tests/golden/fnl_*.npy
// (generated from the vendored header by oracle/make_golden.py).
// ---------------------------------------------------------------------------

static inline void grad3(int idx, float& gx, float& gy, float& gz) {
  // 64-entry table :2529-2547: five repeats of the 12 cube-edge vectors, then 4 extras
  static const signed char base[12][3] = {{0, 1, 1},  {0, -1, 1},  {0, 1, -1},  {0, -1, -1}, {1, 0, 1},  {-1, 0, 1},
                                          {1, 0, -1}, {-1, 0, -1}, {1, 1, 0},   {-1, 1, 0},  {1, -1, 0}, {-1, -1, 0}};
  static const signed char extra[4][3] = {{1, 1, 0}, {0, -1, 1}, {-1, 1, 0}, {0, -1, -1}};
  const signed char* e = idx < 60 ? base[idx % 12] : extra[idx - 60];
  gx = e[0]; gy = e[1]; gz = e[2];
}

static inline float grad_coord(int seed, int xp, int yp, int zp, float xd, float yd, float zd) {
  int hash = seed ^ xp ^ yp ^ zp;
  hash = (int)((unsigned)hash * 0x27d4eb2du);
  hash ^= hash >> 15;
  hash &= 63 << 2;
  float gx, gy, gz;
  grad3(hash >> 2, gx, gy, gz);
  return xd * gx + yd * gy + zd * gz;
}

static inline int fast_round(float f) { return f >= 0 ? (int)(f + 0.5f) : (int)(f - 0.5f); }

static float open_simplex2_3d(int seed, float x, float y, float z) {
  const int PX = 501125321, PY = 1136930381, PZ = 1720413743;
  int i = fast_round(x), j = fast_round(y), k = fast_round(z);
  float x0 = (float)(x - i), y0 = (float)(y - j), z0 = (float)(z - k);
  int xs = (int)(-1.0f - x0) | 1, ys = (int)(-1.0f - y0) | 1, zs = (int)(-1.0f - z0) | 1;
  float ax0 = xs * -x0, ay0 = ys * -y0, az0 = zs * -z0;
  i = (int)((unsigned)i * (unsigned)PX);
  j = (int)((unsigned)j * (unsigned)PY);
  k = (int)((unsigned)k * (unsigned)PZ);
  float value = 0;
  float a = (0.6f - x0 * x0) - (y0 * y0 + z0 * z0);
  for (int l = 0;; l++) {
    if (a > 0) value += (a * a) * (a * a) * grad_coord(seed, i, j, k, x0, y0, z0);
    float b = a + 1;
    int i1 = i, j1 = j, k1 = k;
    float x1 = x0, y1 = y0, z1 = z0;
    if (ax0 >= ay0 && ax0 >= az0) {
      x1 += xs; b -= xs * 2 * x1; i1 = (int)((unsigned)i1 - (unsigned)(xs * PX));
    } else if (ay0 > ax0 && ay0 >= az0) {
      y1 += ys; b -= ys * 2 * y1; j1 = (int)((unsigned)j1 - (unsigned)(ys * PY));
    } else {
      z1 += zs; b -= zs * 2 * z1; k1 = (int)((unsigned)k1 - (unsigned)(zs * PZ));
    }
    if (b > 0) value += (b * b) * (b * b) * grad_coord(seed, i1, j1, k1, x1, y1, z1);
    if (l == 1) break;
    ax0 = 0.5f - ax0; ay0 = 0.5f - ay0; az0 = 0.5f - az0;
    x0 = xs * ax0; y0 = ys * ay0; z0 = zs * az0;
    a += (0.75f - ax0) - (ay0 + az0);
    i = (int)((unsigned)i + (unsigned)((xs >> 1) & PX));
    j = (int)((unsigned)j + (unsigned)((ys >> 1) & PY));
    k = (int)((unsigned)k + (unsigned)((zs >> 1) & PZ));
    xs = -xs; ys = -ys; zs = -zs;
    seed = ~seed;
  }
  return value * 32.69428253173828125f;
}

float orc_noise_sample(float x, float y, float z, float frequency, int octaves, float gain, float lacunarity) {
  x *= frequency; y *= frequency; z *= frequency;
  const float R3 = (float)(2.0 / 3.0);
  const float r = (x + y + z) * R3;
  x = r - x; y = r - y; z = r - z;
  float g = gain < 0 ? -gain : gain, amp = g, ampFractal = 1.0f;
  for (int i = 1; i < octaves; i++) { ampFractal += amp; amp *= g; }
  const float bounding = 1 / ampFractal;
  int seed = 1337;
  float sum = 0;
  amp = bounding;
  for (int i = 0; i < octaves; i++) {
    const float noise = open_simplex2_3d(seed++, x, y, z);
    sum += noise * amp;
    amp *= 1.0f + 0.0f * ((noise + 1) * 0.5f - 1.0f);  // Lerp(1, (n+1)/2, weightedStrength = 0)
    x *= lacunarity; y *= lacunarity; z *= lacunarity;
    amp *= gain;
  }
  return sum;
}

// noise.hpp:42-56: value(i) = GetNoise(pos.x / ext.x, pos.y / ext.y, seed)
void orc_noise(float* out, int nx, int ny, float frequency, int octaves, float gain, float lacunarity, float seed,
               float extx, float exty) {
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y)
      out[(size_t)x * ny + y] = orc_noise_sample(x / extx, y / exty, seed, frequency, octaves, gain, lacunarity);
}

// ---------------------------------------------------------------------------
// One erosion step as composed in SURVEY.md section 3.1 (INFERRED from the kernel
// contracts; the reference's `erode` is not in the tree).  Every pointer is a
// host array in the reference layout.  albedo pointers may be null as a group.
// ---------------------------------------------------------------------------

struct model_t {
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

void orc_erode_step(model_t* m, const param_t* prm, int with_creep) {
  const size_t N = (size_t)m->nx * m->ny;
  std::vector<float> dummy3;
  float *abed = m->albedoBedrock, *asurf = m->albedoSurface, *afl = m->albedoFluxFluvial, *adb = m->albedoFluxDebris;
  const bool albedo = asurf != nullptr;
  if (!albedo) {  // the reference kernels always touch albedo; give them neutral scratch
    dummy3.assign(4 * 3 * N, 0.0f);
    abed = dummy3.data(); asurf = abed + 3 * N; afl = asurf + 3 * N; adb = afl + 3 * N;
  }
  std::memset(m->waterFlux, 0, N * 4);
  std::memset(m->massFlux, 0, N * 4);
  std::memset(m->velocityFlux, 0, N * 8);
  std::memset(m->debrisFlux, 0, N * 4);
  std::memset(m->debrisVelFlux, 0, N * 8);
  std::memset(afl, 0, N * 12);
  std::memset(adb, 0, N * 12);
  // north_star item 4 (EMA of the transported maps): the reference's normalise kernels overwrite,
  // which is lrate == 1 (both examples set it so); for lrate < 1 OUR definition is
  // x <- (1 - lrate) * x_old + lrate * x_normalised, applied before mass transfer reads the maps.
  const bool ema = prm->lrate != 1.0f;
  std::vector<float> o_wh, o_ms, o_mo, o_db, o_dv;
  if (ema) {
    o_wh.assign(m->waterHeight, m->waterHeight + N); o_ms.assign(m->mass, m->mass + N);
    o_mo.assign(m->momentum, m->momentum + 2 * N); o_db.assign(m->debris, m->debris + N);
    o_dv.assign(m->debrisVel, m->debrisVel + 2 * N);
  }
  orc_transport_fluvial(m->layers, m->rainfall, m->waterHeight, m->waterFlux, m->mass, m->massFlux, m->momentum,
                        m->velocityFlux, abed, afl, asurf, m->rngFluvial, m->P, m->nx, m->ny, m->scale, prm, 1, 1);
  orc_transport_debris(m->layers, m->debrisVel, m->debrisVelFlux, m->debris, m->debrisFlux, abed, adb, asurf,
                       m->rngDebris, m->P, m->nx, m->ny, m->scale, prm, 1, 1);
  if (ema) {
    const float lr = prm->lrate, kr = 1.0f - prm->lrate;
    auto blend = [&](float* x, const std::vector<float>& o, size_t n) { for (size_t i = 0; i < n; ++i) x[i] = kr * o[i] + lr * x[i]; };
    blend(m->waterHeight, o_wh, N); blend(m->mass, o_ms, N); blend(m->momentum, o_mo, 2 * N);
    blend(m->debris, o_db, N); blend(m->debrisVel, o_dv, 2 * N);
  }
  std::memset(m->delta, 0, N * 8);
  orc_mass_transfer(m->delta, m->layers, m->uplift, m->waterHeight, m->mass, m->momentum, m->debris, m->debrisVel,
                    abed, afl, adb, albedo ? asurf : nullptr, m->nx, m->ny, m->scale, prm);
  if (with_creep) orc_mass_creep(m->delta, m->layers, m->nx, m->ny, m->scale, prm);
  orc_add(m->layers, m->delta, 2 * (long)N);
  orc_layer_merge(m->height, m->layers, (long)N);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// the gradient family, slope and the Gaussian blur.
// grad.cu:22-206, graph.cu:270-311, filter.cu:24-91.
// ---------------------------------------------------------------------------
extern "C" {

// grad.cu:22-89: central difference, falling back to the backward, then the forward one-sided difference,
// then 0, whenever the previous candidate is NaN (a neighbour outside the map reads as NaN)
void orc_gradient(const float* h, float* out, int nx, int ny, const float* scale) {
  const grid_t g{nx, ny};
  const float nan = std::numeric_limits<float>::quiet_NaN();
  auto at = [&](int x, int y) { return g.oob(x, y) ? nan : h[g.flat(x, y)]; };
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const float c = at(x, y), hn0 = at(x - 1, y), hp0 = at(x + 1, y), h0n = at(x, y - 1), h0p = at(x, y + 1);
      const float gxn = (c - hn0) / scale[0], gyn = (c - h0n) / scale[1];
      const float gxp = (hp0 - c) / scale[0], gyp = (h0p - c) / scale[1];
      float gx = 0.5f * (hp0 - hn0) / scale[0];
      float gy = 0.5f * (h0p - h0n) / scale[1];
      if (gx != gx) gx = gxn;
      if (gx != gx) gx = gxp;
      if (gx != gx) gx = 0.0f;
      if (gy != gy) gy = gyn;
      if (gy != gy) gy = gyp;
      if (gy != gy) gy = 0.0f;
      out[2 * (size_t)g.flat(x, y) + 0] = gx;
      out[2 * (size_t)g.flat(x, y) + 1] = gy;
    }
}

// grad.cu:102-130: length of the per-axis steepest downhill slope; glm::max(a, b) = (a < b) ? b : a
void orc_negslope(const float* h, float* out, int nx, int ny, const float* scale) {
  const grid_t g{nx, ny};
  const float nan = std::numeric_limits<float>::quiet_NaN();
  auto at = [&](int x, int y) { return g.oob(x, y) ? nan : h[g.flat(x, y)]; };
  auto mx = [](float a, float b) { return a < b ? b : a; };
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y) {
      const float c = at(x, y), hn0 = at(x - 1, y), hp0 = at(x + 1, y), h0n = at(x, y - 1), h0p = at(x, y + 1);
      float gx = 0.0f, gy = 0.0f;
      if (hn0 == hn0) gx = mx(gx, (c - hn0) / scale[0]);
      if (hp0 == hp0) gx = mx(gx, (c - hp0) / scale[0]);
      if (h0n == h0n) gy = mx(gy, (c - h0n) / scale[1]);
      if (h0p == h0p) gy = mx(gy, (c - h0p) / scale[1]);
      out[g.flat(x, y)] = sqrtf(gx * gx + gy * gy);  // glm::length
    }
}

// grad.cu:137-176: 9-point Laplacian per component with boundary continuation (a neighbour outside the map
// reads as the centre value); channels = 1 or 2, interleaved
void orc_laplacian(const float* in, float* out, int nx, int ny, int channels, const float* scale) {
  const grid_t g{nx, ny};
  const float hx = (1.0f / scale[0] / scale[0]);
  const float hy = (1.0f / scale[1] / scale[1]);
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int x = 0; x < nx; ++x)
    for (int y = 0; y < ny; ++y)
      for (int ch = 0; ch < channels; ++ch) {
        auto at = [&](int dx, int dy) {
          return g.oob(x + dx, y + dy) ? in[(size_t)g.flat(x, y) * channels + ch]
                                       : in[(size_t)g.flat(x + dx, y + dy) * channels + ch];
        };
        const float v00 = at(0, 0);
        const float vn0 = at(-1, 0), vp0 = at(1, 0), v0n = at(0, -1), v0p = at(0, 1);
        const float vnn = at(-1, -1), vpp = at(1, 1), vpn = at(1, -1), vnp = at(-1, 1);
        const float LH = (vn0 - v00) * hx + (vp0 - v00) * hx + (v0n - v00) * hy + (v0p - v00) * hy;
        const float LD = 0.5f * (vnn - v00) * hx + 0.5f * (vpp - v00) * hx + 0.5f * (vpn - v00) * hy + 0.5f * (vnp - v00) * hy;
        out[(size_t)g.flat(x, y) * channels + ch] = 0.5f * LH + 0.5f * LD;
      }
}

// graph.cu:270-296: slope along the flow edge; 0 where the cell has no receiver (or points to itself)
void orc_slope(const float* h, const int* flow, float* out, int nx, int ny, const float* scale) {
  const grid_t g{nx, ny};
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (int n = 0; n < nx * ny; ++n) {
    const int next = flow[n];
    if (next < 0 || next == n) { out[n] = 0.0f; continue; }
    const float ix = (float)(n / ny), iy = (float)(n % ny);
    const float px = (float)(next / ny), py = (float)(next % ny);
    const float dx = scale[0] * (px - ix), dy = scale[1] * (py - iy);
    out[n] = (h[next] - h[n]) / sqrtf(dx * dx + dy * dy);
  }
  (void)g;
}

// filter.cu:24-91: separable 33-tap Gaussian, first along dim 0 (inout -> tmp), then along dim 1
// (tmp -> inout): the argument is overwritten and is the result.  Indices clamp at the map edge, the weights
// are not renormalised there.  __expf on the GPU, expf here.
void orc_gaussian_blur(float* inout, float* tmp, int nx, int ny, int channels, float sigma) {
  const grid_t g{nx, ny};
  const int kwindow = 16;
  auto pass = [&](float* dst, const float* src, bool xdir) {
#pragma omp parallel for num_threads(nthreads()) schedule(static)
    for (int x = 0; x < nx; ++x)
      for (int y = 0; y < ny; ++y)
        for (int ch = 0; ch < channels; ++ch) {
          float val = 0.0f;
          for (int k = -kwindow; k <= kwindow; ++k) {
            int px = x + (xdir ? k : 0), py = y + (xdir ? 0 : k);
            if (px < 0) px = 0;
            if (py < 0) py = 0;
            if (px > nx - 1) px = nx - 1;
            if (py > ny - 1) py = ny - 1;
            const float Z = sqrtf(2.0f * 3.14159265f) * sigma;
            const float kernel = expf(-0.5f * (k / sigma) * (k / sigma)) / Z;
            val += src[(size_t)g.flat(px, py) * channels + ch] * kernel;
          }
          dst[(size_t)g.flat(x, y) * channels + ch] = val;
        }
  };
  pass(tmp, inout, true);
  pass(inout, tmp, false);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// SURVEY.md section 8(f) row 4 (first half): the albedo generators, erosion.cu:759-919.
// All per-cell, albedo maps are (nx, ny, 3) interleaved.
// ---------------------------------------------------------------------------
extern "C" {

// erosion.cu:759-777
void orc_albedo_layer(float* albedo, const float* albedoBedrock, const float* albedoSediment, const float* layers,
                      long n, float scaleSediment, const float* shiftSediment) {
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) {
    const float blend = 1.0f / (1.0f + scaleSediment * layers[2 * i + 1]);
    for (int c = 0; c < 3; ++c) {
      const float colorSediment = fminf(albedoSediment[3 * i + c] + shiftSediment[c], 1.0f);  // __mmin(1, .)
      albedo[3 * i + c] = blend * albedoBedrock[3 * i + c] + (1.0f - blend) * colorSediment;
    }
  }
}

// erosion.cu:793-826: alternating strata by the depth below the total uplift displacement
void orc_albedo_stratum(float* albedoBedrock, const float* uplift, const float* layers, long n, const float* scale,
                        const param_t* param, const float* colorA, const float* colorB, float age, float freq) {
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) {
    const float shift = age * param->uplift * uplift[i];
    const float depth = fmaxf(shift - layers[2 * i] * scale[2], 0.0f);
    const int index = (int)floorf(depth / freq);
    const float* col = (index % 2 == 0) ? colorA : colorB;
    for (int c = 0; c < 3; ++c) albedoBedrock[3 * i + c] = col[c];
  }
}

// erosion.cu:858-876; __expf on the GPU, expf here
void orc_albedo_discharge(float* albedo, const float* discharge, long n, const float* colorDischarge,
                          float extinction, float scale) {
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < n; ++i) {
    const float value = fmaxf(0.0f, discharge[i]);
    const float blend = scale * (1.0f - expf(-extinction * value));
    for (int c = 0; c < 3; ++c) albedo[3 * i + c] = blend * colorDischarge[c] + (1.0f - blend) * albedo[3 * i + c];
  }
}

}  // extern "C"

// ---------------------------------------------------------------------------
// soil::solve_uniform, path.cu:51-219 with the bilinear flow sampling of
// sample.hpp:72-186 (the view overload, row-major).  A uniform-start Monte-Carlo estimator of a linear
// conservation law: each sample follows the flow field and credits every cell it enters with its source rate,
// attenuated by the decay along the way; then everything is normalised together.
// Conversions float -> int truncate; a NaN position converts to 0 like cvt.rzi on the GPU.
// ---------------------------------------------------------------------------
namespace {

inline int trunc_i(float f) { return f != f ? 0 : (int)f; }

// sample_t<vec2,2,1>::gather(view, res, pos).val(), sample.hpp:151-186 and :86-88
inline v2 flow_at(const v2* flow, int nx, int ny, v2 pos) {
  const float nan = std::numeric_limits<float>::quiet_NaN();
  if (pos.x < 0 || pos.y < 0 || pos.x > nx - 1 || pos.y > ny - 1) return v2{nan, nan};
  const int px = trunc_i(pos.x), py = trunc_i(pos.y);
  float wx = pos.x - floorf(pos.x), wy = pos.y - floorf(pos.y);
  int i00 = px * ny + py, i01 = px * ny + (py + 1), i10 = (px + 1) * ny + py, i11 = (px + 1) * ny + (py + 1);
  if (pos.x + 1 > nx - 1) { wx = 0; i10 = 0; i11 = 0; }
  if (pos.y + 1 > ny - 1) { wy = 0; i01 = 0; i11 = 0; }
  const v2 h00 = flow[i00], h01 = flow[i01], h10 = flow[i10], h11 = flow[i11];
  // M * (1, t) = (1 + (-1) t, 0 + t): the weights as the matrix product forms them
  const float ay = 1.0f * 1.0f + -1.0f * wy, by = 0.0f * 1.0f + 1.0f * wy;
  const v2 l0 = {ay * h00.x + by * h01.x, ay * h00.y + by * h01.y};
  const v2 l1 = {ay * h10.x + by * h11.x, ay * h10.y + by * h11.y};
  const float ax = 1.0f * 1.0f + -1.0f * wx, bx = 0.0f * 1.0f + 1.0f * wx;
  return v2{ax * l0.x + bx * l1.x, ax * l0.y + bx * l1.y};
}

}  // namespace

extern "C" {

// flow (nx,ny,2), source (nx,ny,K), decay (nx,ny), rng: P XORWOW states (advanced by two draws each),
// flux (nx,ny,K) is overwritten with the normalised estimate.  K = 1 or 2.
void orc_solve_uniform(const float* flow_, const float* source, const float* decay, void* rng_, long P, float* flux,
                       int nx, int ny, int K, const float* scale, long count) {
  const v2* flow = (const v2*)flow_;
  curandStateXORWOW_t* rng = (curandStateXORWOW_t*)rng_;
  const size_t N = (size_t)nx * ny;
  std::memset(flux, 0, N * K * sizeof(float));
  const float epsilon = 1E-16f;
  const float maxstep = (float)(nx + ny);
  const float L = sqrtf(scale[0] * scale[0] + scale[1] * scale[1]);
  const float A = scale[0] * scale[1];
  const float Pr = 1.0f / (A * (float)N);
  for (long n = 0; n < P; ++n) {  // serial: the flux sums are order dependent only in the last bits
    float att = 1.0f;
    v2 pos;
    pos.x = curand_uniform(&rng[n]) * float(nx);
    pos.y = curand_uniform(&rng[n]) * float(ny);
    auto oob = [&](v2 p) { const int x = trunc_i(p.x), y = trunc_i(p.y); return x < 0 || y < 0 || x >= nx || y >= ny; };
    if (oob(pos)) continue;  // a draw of exactly 1.0 starts outside (App. B-12); the reference reads out of bounds there
    int ind = trunc_i(pos.x) * ny + trunc_i(pos.y);
    float S[2] = {0.0f, 0.0f};
    for (int c = 0; c < K; ++c) S[c] = source[(size_t)ind * K + c] / Pr;
    const float lenS = K == 1 ? sqrtf(S[0] * S[0]) : sqrtf(S[0] * S[0] + S[1] * S[1]);
    if (lenS < epsilon) continue;
    int step = 0;
    while (!oob(pos) && epsilon < fabsf(att) && (float)(++step) < maxstep) {
      const int nind = trunc_i(pos.x) * ny + trunc_i(pos.y);
      if (nind != ind) {
        ind = nind;
        for (int c = 0; c < K; ++c) flux[(size_t)ind * K + c] += S[c] * att;
      }
      const v2 v = flow_at(flow, nx, ny, pos);
      const float v_len = sqrtf(v.x * v.x + v.y * v.y);
      if (v_len < epsilon) break;
      const v2 v_norm = {v.x / v_len, v.y / v_len};
      const float st = stepsize(pos, v_norm);  // path.cu:26-49, the same voxel step as erosion_map.cu:56-78
      pos.x += st * v_norm.x;
      pos.y += st * v_norm.y;
      const float dlambda = st * L / v_len;
      att *= expf(-dlambda * decay[ind]);  // __expf on the GPU
    }
  }
  // path.cu:141-167
#pragma omp parallel for num_threads(nthreads()) schedule(static)
  for (long i = 0; i < (long)N; ++i) {
    const v2 v = flow[i];
    const float norm = fabsf(v.x * scale[1]) + fabsf(v.y * scale[0]);
    for (int c = 0; c < K; ++c)
      flux[(size_t)i * K + c] = (source[(size_t)i * K + c] * A + flux[(size_t)i * K + c] / float(count)) / norm;
  }
}

}  // extern "C"
