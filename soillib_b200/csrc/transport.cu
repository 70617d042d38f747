// transport.cu - stochastic particle descent: fluvial and debris-flow transport.
//
// Reference: source/soillib/model/path/erosion.cu:29-141 (fluvial particle),
// :245-351 (debris particle).  One particle per thread; a particle is a serial
// chain of <= maxage cell crossings, each a dependent gather of the 5-point
// height stencil + velocity + water height followed by fire-and-forget
// reductions into the flux maps.  The trajectory arithmetic is the reference's,
// operand for operand (no FMA contraction), so a particle visits exactly the
// cells the CPU oracle's particle visits.
//
// Two memory layouts, selected by a policy type:
//  * ref_io   - the reference's tensors as they are (per-op entry points):
//               heights from the interleaved `layers`, one RED per flux float.
//  * fused_io - the fused step's private layout: heights from the merged
//               `height` map (the dense pass writes bedrock + sediment there, the
//               same float add the reference performs per access), and ONE packed
//               flux record per cell and particle system, updated with 16-byte
//               vector reductions (red.global.add.v4.f32): a crossing dirties one
//               32-byte sector instead of five, and sets the cell's tag bit for the
//               dense pass.
#include "common.cuh"
#include "erosion_math.cuh"
#include <curand_kernel.h>
#include <cstdlib>

namespace sb {
namespace {

__device__ __forceinline__ bool oobf(float px, float py, int nx, int ny) {  // erosion_map.cu:29-40
  if (px < 0) return true;
  if (py < 0) return true;
  if (px >= nx) return true;
  if (py >= ny) return true;
  return false;
}

// shape.flatten(ivec2(pos)): truncation toward zero, NaN -> 0 (cvt.rzi), App. A-5 / B-4
__device__ __forceinline__ int flattenf(float px, float py, int ny) { return (int)px * ny + (int)py; }

__device__ __forceinline__ void red_add2(float2* addr, float vx, float vy) {
  atomicAdd(addr, make_float2(vx, vy));  // red.global.add.v2.f32 (sm_90+)
}
__device__ __forceinline__ void red_add4(float* addr, float a, float b, float c, float d) {
  atomicAdd(reinterpret_cast<float4*>(addr), make_float4(a, b, c, d));  // red.global.add.v4.f32 (sm_90+)
}

// Reductions that may land in ANOTHER GPU's memory (peer-mapped strips): system scope, because the owner and other
// peers update the same records concurrently and .gpu scope does not order against threads of another GPU; written
// as `red` (no return value) on a global address, so the particle does not wait for the round trip over NVLink -
// atomicAdd() on a generic pointer compiles to an ATOM that returns a predicate plus a shared-memory fallback loop.
__device__ __forceinline__ void red_add_sys(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add4_sys(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(p)), "f"(a),
               "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_or_sys(unsigned* p, unsigned v) {
  asm volatile("red.relaxed.sys.global.or.b32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "r"(v) : "memory");
}

// ---- layout policies ---------------------------------------------------------------

// where the fused step's maps of ONE rank live (an array of these, one per rank, sits in device memory when
// the strips of all ranks are reachable through peer mappings); 64 bytes, mirrors soilb200_peer_t
struct peer_dev {
  int own_lo, own_hi, row0, rows;
  const float* hmap;
  const float* wh;
  const float2* mom;
  const float2* dmom;
  float* flux;
  unsigned char* touched;
  const float4* view;  // packed particle view of the rank (both particle systems), or null
  const void* pad;
};
static_assert(sizeof(peer_dev) == 80, "peer_dev");

// ---- packed particle view -----------------------------------------------------------------------------------
// What a particle gathers at a crossing - the height of the cell and of its four neighbours, and the velocity and
// water height of the cell - comes from seven scattered 4/8-byte loads out of three row-major maps: five 64-byte DRAM
// requests per crossing, in a kernel that is bound by the rate of random DRAM requests (profiles/r02_*).  In view
// mode the dense pass writes, per particle system, ONE 16-byte record per cell instead - fluvial {height, water height,
// velocity.x, velocity.y}, debris {height, velocity.x, velocity.y, debris mass} - in 2 x 2-cell quads (64 bytes = one
// DRAM request): the cell's record brings its height, velocity and water height in one load, one x- and one
// y-neighbour sit in the same quad, and the step to the next cell lands in a quad the last crossing fetched.
// (a view of a map with nx * ny < 2^31 cells has fewer than 2^31 + 2 (nx + ny) records: 32-bit indices)
__device__ __forceinline__ unsigned quad_index(int lx, int y, int nyq) {
  return (((unsigned)(lx >> 1) * (unsigned)nyq + (unsigned)(y >> 1)) << 2) + (unsigned)(((lx & 1) << 1) + (y & 1));
}

struct ref_io {
  const float2* layers;
  // fluvial: water, mass, velocity(2), albedo(3); debris: mass, velocity(2), albedo(3)
  float *waterFlux, *massFlux;
  float2* velocityFlux;
  float* albedoFlux;  // may be null
  const float* wh;    // water height (fluvial only)
  const float2* vel;  // the particle system's velocity map
  size_t off;         // always 0: the arrays are the whole map
  __device__ __forceinline__ bool relocate(int) { return true; }
  __device__ __forceinline__ float height(size_t i) const {
    const float2 l = __ldg(layers + i);
    return l.x + l.y;  // erosion_map.cu:99-105
  }
  __device__ __forceinline__ bool albedo() const { return albedoFlux != nullptr; }
  __device__ __forceinline__ void deposit_fluvial(int ind, float w, float m, float vx, float vy, float ar, float ag,
                                                  float ab) const {
    atomicAdd(waterFlux + ind, w);
    atomicAdd(massFlux + ind, m);
    red_add2(velocityFlux + ind, vx, vy);
    if (albedoFlux) {
      atomicAdd(albedoFlux + 3 * (size_t)ind + 0, ar);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 1, ag);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 2, ab);
    }
  }
  __device__ __forceinline__ void deposit_debris(int ind, float m, float vx, float vy, float ar, float ag,
                                                 float ab) const {
    atomicAdd(massFlux + ind, m);
    red_add2(velocityFlux + ind, vx, vy);
    if (albedoFlux) {
      atomicAdd(albedoFlux + 3 * (size_t)ind + 0, ar);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 1, ag);
      atomicAdd(albedoFlux + 3 * (size_t)ind + 2, ab);
    }
  }
};

// record of R = 4 (no albedo) or 8 floats per cell:
//   fluvial {water, mass, vel.x, vel.y | r, g, b, 0}   debris {mass, vel.x, vel.y, 0 | r, g, b, 0}
struct fused_io {
  // gathers: the arrays they currently go to and the first global cell those hold
  const float* hmap;
  const float* wh;
  const float2* vel;
  size_t roff;
  bool far;  // the gathers currently go to another GPU's arrays (over NVLink)
  // deposits: records and tags of the rank that owns the particle's current rows (its own to begin with)
  float* rec;  // this particle system's records
  int R;
  unsigned char* touched;
  size_t off;  // first global cell of the deposit arrays: row0 * ny
  int own_lo, own_hi;
  bool remote;  // the deposit arrays belong to another GPU
  // this rank's own arrays hold the global rows [win_lo, win_hi): its rows plus read-only halo rows that the
  // caller refreshes every step.  While a particle's 5-point stencil stays inside that window its gathers stay
  // local even in a neighbour's rows - only the (fire-and-forget) deposits cross NVLink, and no load latency
  // of a remote GPU ends up on the particle's dependent chain.
  const float* my_hmap;
  const float* my_wh;
  const float2* my_vel;
  size_t my_off;
  int win_lo, win_hi, nx_g;
  // view mode (null: the row-major maps above): this particle system's records of the rank the gathers currently go
  // to, the global row of its local row 0, quads per row pair; the rank's own view
  const float4* view;
  int vrow0, nyq;
  const float4* my_view;
  int my_vrow0;
  // every rank's maps, when peer-mapped (null: particles that leave the owned rows are exported instead)
  const peer_dev* peers;
  int n_peers, self, kind, ny;

  __device__ __forceinline__ bool in_window(int gx) const {
    return max(gx - 1, 0) >= win_lo && min(gx + 1, nx_g - 1) < win_hi;
  }
  __device__ __forceinline__ bool relocate(int gx) {
    if (!peers) return gx >= own_lo && gx < own_hi;   // one rank's rows: nothing to look up
    // same owner as at the last crossing and the same choice of where to gather from: nothing to look up either
    if (gx >= own_lo && gx < own_hi && far == (remote && !in_window(gx))) return true;
    for (int r = 0; r < n_peers; ++r) {
      const int lo = __ldg(&peers[r].own_lo), hi = __ldg(&peers[r].own_hi);
      if (gx < lo || gx >= hi) continue;
      const peer_dev p = peers[r];
      own_lo = lo; own_hi = hi;
      off = (size_t)p.row0 * ny;
      touched = p.touched;
      rec = p.flux + (kind ? R : 0);
      remote = r != self;
      if (in_window(gx)) {
        hmap = my_hmap; wh = my_wh; vel = my_vel; roff = my_off; far = false;
        view = my_view; vrow0 = my_vrow0;
      } else {  // beyond the halo rows (deeper than the caller provisioned): gather from the owner
        hmap = p.hmap; wh = p.wh; vel = kind ? p.dmom : p.mom; roff = off; far = remote;
        if (my_view) { view = p.view + (kind ? (size_t)((p.rows + 1) >> 1) * nyq * 4 : 0); vrow0 = p.row0; }
      }
      return true;
    }
    return false;
  }
  __device__ __forceinline__ float height(size_t i) const { return __ldg(hmap + i); }
  __device__ __forceinline__ bool albedo() const { return R == 8; }
  __device__ __forceinline__ void add4(float* r, float a, float b, float c, float d) const {
    if (!peers) { red_add4(r, a, b, c, d); return; }          // one GPU: device scope
    if (!remote) { red_add4_sys(r, a, b, c, d); return; }     // own records that peers also update
    red_add_sys(r, a); red_add_sys(r + 1, b); red_add_sys(r + 2, c);  // peer memory: 4-byte reductions over NVLink
    if (d != 0.0f) red_add_sys(r + 3, d);
  }
  __device__ __forceinline__ void tag(int ind) const {  // one tag bit per cell
    unsigned* w = reinterpret_cast<unsigned*>(touched) + (ind >> 5);
    if (!peers) atomicOr(w, 1u << (ind & 31));
    else red_or_sys(w, 1u << (ind & 31));
  }
  __device__ __forceinline__ void deposit_fluvial(int ind, float w, float m, float vx, float vy, float ar, float ag,
                                                  float ab) const {
    float* r = rec + (size_t)2 * R * ind;  // {fluvial record, debris record} per cell
    add4(r, w, m, vx, vy);
    if (R == 8) add4(r + 4, ar, ag, ab, 0.0f);
    tag(ind);
  }
  __device__ __forceinline__ void deposit_debris(int ind, float m, float vx, float vy, float ar, float ag,
                                                 float ab) const {
    float* r = rec + (size_t)2 * R * ind;  // {fluvial record, debris record} per cell
    add4(r, m, vx, vy, 0.0f);
    if (R == 8) add4(r + 4, ar, ag, ab, 0.0f);
    tag(ind);
  }
};

// ---- row strips (multi-GPU) ------------------------------------------------------------------
// Positions and cell indices are always GLOBAL.  The local arrays hold the global rows
// [row0, row0 + allocated rows); this rank owns [own_lo, own_hi) and keeps one halo row on each side
// that has a neighbour.  A particle whose current cell is not owned is written, whole, to the outbox of
// the neighbour it moved towards and continues there (k_transport_resume) with the same arithmetic, so
// the union over ranks is the single-GPU particle set.  Single domain: row0 = 0, own = [0, nx_g).
struct strip_t {
  int row0, own_lo, own_hi, nx_g;
  float *outbox_lo, *outbox_hi;  // towards lower / higher rows; 16 floats per record
  int* counts;       // [2] records appended (may exceed capacity: the host checks)
  int capacity;
};

struct pstate {  // one particle in flight, 64 bytes
  float px, py, sx, sy;
  float att0, att1, att2;  // fluvial: water, mass, velocity   debris: debris, velocity, -
  float s0, s1, s2, s3, s4, s5, s6;  // fluvial: w, m, vx, vy, ar, ag, ab   debris: d, vx, vy, ar, ag, ab, -
  int ind;   // global flat index of the last credited cell
  int iter;  // crossings so far | kind << 30 (0 fluvial, 1 debris)
};
static_assert(sizeof(pstate) == 64, "pstate");

__device__ __forceinline__ void export_particle(const strip_t& sp, const pstate& st, int dir) {
  const int slot = atomicAdd(sp.counts + dir, 1);
  if (slot < sp.capacity) {
    float4* dst = reinterpret_cast<float4*>((dir ? sp.outbox_hi : sp.outbox_lo) + 16 * (size_t)slot);
    dst[0] = make_float4(st.px, st.py, st.sx, st.sy);  // built from registers: the state never lives in local memory
    dst[1] = make_float4(st.att0, st.att1, st.att2, st.s0);
    dst[2] = make_float4(st.s1, st.s2, st.s3, st.s4);
    dst[3] = make_float4(st.s5, st.s6, __int_as_float(st.ind), __int_as_float(st.iter));
  }
}

// the five heights __glocal (erosion_map.cu:107-159) reads at global cell (gx,gy); NaN = outside the map
struct h5 {
  float c, n0, p0, zn, zp;
};
// everything a crossing gathers
struct cell_in {
  h5 w;
  float2 vel;
  float wh;
};
template <class IO>
__device__ __forceinline__ h5 load5(const IO& io, int gx, int gy, size_t off, int nx_g, int ny) {
  const size_t i = (size_t)gx * ny + gy - off;
  h5 w;
  w.n0 = w.p0 = w.zn = w.zp = CUDART_NAN_F;
  w.c = io.height(i);
  if (gx > 0) w.n0 = io.height(i - ny);
  if (gx + 1 < nx_g) w.p0 = io.height(i + ny);
  if (gy > 0) w.zn = io.height(i - 1);
  if (gy + 1 < ny) w.zp = io.height(i + 1);
  return w;
}
// the gathers of one crossing at global cell (gx, gy) through the layout policy's current pointers
__device__ __forceinline__ cell_in gather(const ref_io& io, int gx, int gy, int nx_g, int ny, bool want_wh) {
  cell_in c;
  const size_t i = (size_t)gx * ny + gy;
  c.w = load5(io, gx, gy, 0, nx_g, ny);
  c.vel = __ldg(io.vel + i);
  c.wh = want_wh ? __ldg(io.wh + i) : 0.0f;
  return c;
}
__device__ __forceinline__ cell_in gather(const fused_io& io, int gx, int gy, int nx_g, int ny, bool want_wh) {
  cell_in c;
  if (io.view) {
    const int lx = gx - io.vrow0, nyq = io.nyq;
    const unsigned q = quad_index(lx, gy, nyq);
    const float4 r = __ldg(io.view + q);
    c.w.n0 = c.w.p0 = c.w.zn = c.w.zp = CUDART_NAN_F;
    c.w.c = r.x;
    // the neighbours as offsets from the cell's record: inside its quad (+-1, +-2) or in the quad beside it
    const unsigned row = 4u * (unsigned)nyq;
    const unsigned qxm = (lx & 1) ? q - 2u : q + 2u - row, qxp = (lx & 1) ? q - 2u + row : q + 2u;
    const unsigned qym = (gy & 1) ? q - 1u : q - 3u, qyp = (gy & 1) ? q + 3u : q + 1u;
    if (gx > 0) c.w.n0 = __ldg(reinterpret_cast<const float*>(io.view + qxm));
    if (gx + 1 < nx_g) c.w.p0 = __ldg(reinterpret_cast<const float*>(io.view + qxp));
    if (gy > 0) c.w.zn = __ldg(reinterpret_cast<const float*>(io.view + qym));
    if (gy + 1 < ny) c.w.zp = __ldg(reinterpret_cast<const float*>(io.view + qyp));
    // fluvial {h, wh, vx, vy}, debris {h, vx, vy, debris mass}
    c.vel = io.kind ? make_float2(r.y, r.z) : make_float2(r.z, r.w);
    c.wh = r.y;
    return c;
  }
  const size_t i = (size_t)gx * ny + gy - io.roff;
  c.w = load5(io, gx, gy, io.roff, nx_g, ny);
  c.vel = __ldg(io.vel + i);
  c.wh = want_wh ? __ldg(io.wh + i) : 0.0f;
  return c;
}

__device__ __forceinline__ f2 grad_of(const h5& w, const scale3 s, float exitSlope) {
  return glocal_h(w.c, w.n0, w.p0, w.zn, w.zp, s, exitSlope);
}

struct particle_ctx {
  curandStateXORWOW_t* rng;
  long long P;
  int ny;
  scale3 scale;
  const float* albedoSource;  // (rows,ny,3) or null
  strip_t sp;
  // strips: start positions of the particles that start in this rank's rows, drawn and compacted by k_spawn
  // (null: every thread draws its own start from rng[n])
  const float2* spawn_pos;
  const int* spawn_count;
  __device__ __forceinline__ bool start(long long n, int nx, float& px, float& py) const {
    if (spawn_pos) {
      if (n >= (long long)__ldg(spawn_count)) return false;
      const float2 p = __ldg(spawn_pos + n);
      px = p.x; py = p.y;
      return true;
    }
    curandStateXORWOW_t rs = rng[n];
    px = 0.5f + curand_uniform(&rs) * float(nx - 1);  // erosion.cu:56-59
    py = 0.5f + curand_uniform(&rs) * float(ny - 1);
    rng[n] = rs;
    return true;
  }
  __device__ __forceinline__ size_t off() const { return (size_t)sp.row0 * ny; }
  __device__ __forceinline__ bool mine(int gx) const { return gx >= sp.own_lo && gx < sp.own_hi; }
};

// ---- fluvial particle, erosion.cu:46-139 ---------------------------------------------

// start position, initial speed and the source terms fixed at the start cell (erosion.cu:46-96)
template <class IO>
__device__ __forceinline__ bool fluvial_init(const long long n, const IO& io, const particle_ctx& c,
                                             const float* __restrict__ waterSource, const soilb200_param_t& param,
                                             pstate& st) {
  const int nx = c.sp.nx_g, ny = c.ny;
  const scale3 scale = c.scale;
  const float A = scale.x * scale.y;
  const float Lx = scale.x, Ly = scale.y;
  const float Pp = 1.0f / float(A * (float)((size_t)nx * (size_t)ny));
  const float Q = 1.0f / (Pp * (float)(size_t)c.P);
  const float eps = 1E-12f;

  float px, py;
  if (!c.start(n, nx, px, py)) return false;
  const int gx = (int)px, gy = (int)py;
  if (!c.mine(gx)) return false;
  const int ind = flattenf(px, py, ny);
  const size_t li = (size_t)ind - c.off();

  const float rho_w = param.densityWater;
  const float nu = param.viscosityWater;
  const float g = param.gravity;
  const float ks = param.suspensionRateFluvial / 64.0f;
  const float fD = param.frictionFactor / 8.0f;
  const float alpha = param.fluvialExponent;
  const float R = param.rainfall;

  const cell_in ci = gather(io, gx, gy, nx, ny, false);
  const float2 vel = ci.vel;
  const f2 grad = grad_of(ci.w, scale, param.exitSlope);
  float sx = -(g * grad.x) + nu * vel.x + param.force[0];
  float sy = -(g * grad.y) + nu * vel.y + param.force[1];
  {
    const float d = sqrtf(len2(Lx * sx, Ly * sy));
    sx = sx / d;
    sy = sy / d;
  }
  if (len2(sx, sy) < eps) return false;

  const float v = len2(vel.x, vel.y);
  const float shear = 0.125f * fD * rho_w * v * v;
  const float power = __powf(shear * len2(grad.x, grad.y), alpha);

  const float source_m = Q * ks * power;
  st.px = px; st.py = py; st.sx = sx; st.sy = sy;
  st.att0 = 1.0f; st.att1 = 1.0f; st.att2 = 1.0f;
  st.s0 = Q * R * __ldg(waterSource + li);
  st.s1 = source_m;
  st.s2 = Q * (-(g * grad.x) + nu * vel.x);
  st.s3 = Q * (-(g * grad.y) + nu * vel.y);
  st.s4 = st.s5 = st.s6 = 0.0f;
  if (io.albedo()) {
    st.s4 = source_m * __ldg(c.albedoSource + 3 * li + 0);
    st.s5 = source_m * __ldg(c.albedoSource + 3 * li + 1);
    st.s6 = source_m * __ldg(c.albedoSource + 3 * li + 2);
  }
  st.ind = ind;
  st.iter = 0;
  return true;
}

// the trajectory loop (erosion.cu:98-139), resumable
template <class IO>
__device__ __forceinline__ void fluvial_run(pstate st, IO io, const particle_ctx& c, const soilb200_param_t& param) {
  const int nx = c.sp.nx_g, ny = c.ny;
  const scale3 scale = c.scale;
  const float eps = 1E-12f;
  const float tau = param.bedShearWater;
  const float nu = param.viscosityWater;
  const float g = param.gravity;
  const float kd = param.depositionRateFluvial * 1.33f;
  const float fD = param.frictionFactor / 8.0f;
  const float lenL = len2(scale.x, scale.y);

  float px = st.px, py = st.py, sx = st.sx, sy = st.sy;
  float att_w = st.att0, att_m = st.att1, att_v = st.att2;
  int ind = st.ind;
  int iter = st.iter;

  // The gathers of the NEXT crossing are issued as soon as the next position is known (right after the step
  // size), so their DRAM latency overlaps the rest of this crossing's arithmetic: `have` says the window
  // (hw, vl, wh) already holds the current cell.  Same loads, same values, same order of arithmetic.
  bool have = false;
  h5 hw;
  float2 vl;
  float wh;
  hw.c = hw.n0 = hw.p0 = hw.zn = hw.zp = 0.0f; vl = make_float2(0.0f, 0.0f); wh = 0.0f;
  while (true) {
    if (oobf(px, py, nx, ny)) break;
    if (!((unsigned long long)(++iter) < param.maxage)) break;
    const int gx = (int)px, gy = (int)py;
    if (!have) {
      if (!io.relocate(gx)) {  // a neighbour's rows that are not peer-mapped: hand the particle over, this crossing included
        st.px = px; st.py = py; st.sx = sx; st.sy = sy;
        st.att0 = att_w; st.att1 = att_m; st.att2 = att_v;
        st.ind = ind; st.iter = iter - 1;
        export_particle(c.sp, st, gx < c.sp.own_lo ? 0 : 1);
        return;
      }
      const cell_in ci = gather(io, gx, gy, nx, ny, true);
      hw = ci.w; vl = ci.vel; wh = ci.wh;
    }
    const int nind = flattenf(px, py, ny);
    const size_t li = (size_t)nind - io.off;
    if (nind != ind) {
      ind = nind;
      io.deposit_fluvial((int)li, att_w * st.s0, att_m * st.s1, att_v * st.s2, att_v * st.s3, att_m * st.s4,
                         att_m * st.s5, att_m * st.s6);
    }

    const float v_norm = len2(sx, sy);
    const float ux = sx / v_norm, uy = sy / v_norm;   // (zero components are rare here, unlike in the debris loop)
    const float v_step = stepsize(px, py, ux, uy);
    const float dL = v_step * lenL;
    const float ds = div_zn(dL, v_norm);
    if (v_norm < eps) break;

    // next position (erosion.cu:137) - known here; start its gathers now
    const float npx = px + v_step * ux, npy = py + v_step * uy;
    have = false;
    h5 nhw = hw;
    float2 nvl = vl;
    float nwh = wh;
    if (!oobf(npx, npy, nx, ny) && (unsigned long long)(iter + 1) < param.maxage) {
      const int ngx = (int)npx, ngy = (int)npy;
      if (io.relocate(ngx)) {
        const cell_in ci = gather(io, ngx, ngy, nx, ny, true);
        nhw = ci.w; nvl = ci.vel; nwh = ci.wh;
        have = true;
      }
    }

    const f2 grad = grad_of(hw, scale, param.exitSlope);
    const float ax = -(g * grad.x) + nu * vl.x + param.force[0];
    const float ay = -(g * grad.y) + nu * vl.y + param.force[1];
    const float c0 = 1.0f / (1.0f + dL * (tau + nu));
    const float c1 = div_zn(dL, 1.0f + dL * (tau + nu));
    sx = c0 * sx + c1 * ax;
    sy = c0 * sy + c1 * ay;

    const float decay_m = kd;
    const float decay_w = param.evapRate;
    const float decay_v = 0.125f * fD / (eps + wh);
    att_m = att_m * __expf(-ds * decay_m);
    att_w = att_w * __expf(-ds * decay_w);
    att_v = att_v * __expf(-dL * decay_v);
    px = npx;
    py = npy;
    hw = nhw; vl = nvl; wh = nwh;
  }
}

// ---- debris particle, erosion.cu:259-349 -----------------------------------------------

template <class IO>
__device__ __forceinline__ bool debris_init(const long long n, const IO& io, const particle_ctx& c,
                                            const soilb200_param_t& param, pstate& st) {
  const int nx = c.sp.nx_g, ny = c.ny;
  const scale3 scale = c.scale;
  const float A = scale.x * scale.y;
  const float Lx = scale.x, Ly = scale.y;
  const float Pp = 1.0f / float(A * (float)((size_t)nx * (size_t)ny));
  const float Q = 1.0f / (Pp * (float)(size_t)c.P);
  const float eps = 1E-12f;

  float px, py;
  if (!c.start(n, nx, px, py)) return false;
  const int gx = (int)px, gy = (int)py;
  if (!c.mine(gx)) return false;
  const int ind = flattenf(px, py, ny);
  const size_t li = (size_t)ind - c.off();

  const float theta = param.critSlopeBedrock;
  const float nu = param.viscosityDebris;
  const float g = param.gravity;
  const float kl = param.landslideRateDebris;

  const cell_in ci = gather(io, gx, gy, nx, ny, false);
  const float2 vel = ci.vel;
  const f2 grad = grad_of(ci.w, scale, param.exitSlope);
  float sx = -(g * grad.x) + nu * vel.x;
  float sy = -(g * grad.y) + nu * vel.y;
  {
    const float d = sqrtf(len2(Lx * sx, Ly * sy));
    sx = sx / d;
    sy = sy / d;
  }
  if (len2(sx, sy) < eps) return false;

  const float excessSlope0 = (len2(grad.x, grad.y) - theta);
  const float suspend = fmaxf(0.0f, kl * excessSlope0);
  const float source_d = Q * suspend;
  st.px = px; st.py = py; st.sx = sx; st.sy = sy;
  st.att0 = 1.0f; st.att1 = 1.0f; st.att2 = 0.0f;
  st.s0 = source_d;
  st.s1 = Q * (-g * grad.x + nu * vel.x);
  st.s2 = Q * (-g * grad.y + nu * vel.y);
  st.s3 = st.s4 = st.s5 = st.s6 = 0.0f;
  if (io.albedo()) {
    st.s3 = source_d * __ldg(c.albedoSource + 3 * li + 0);
    st.s4 = source_d * __ldg(c.albedoSource + 3 * li + 1);
    st.s5 = source_d * __ldg(c.albedoSource + 3 * li + 2);
  }
  st.ind = ind;
  st.iter = 1 << 30;
  return true;
}

template <class IO>
__device__ __forceinline__ void debris_run(pstate st, IO io, const particle_ctx& c, const soilb200_param_t& param) {
  const int nx = c.sp.nx_g, ny = c.ny;
  const scale3 scale = c.scale;
  const float eps = 1E-12f;
  const float theta = param.critSlopeBedrock;
  const float nu = param.viscosityDebris;
  const float tau = param.bedShearDebris;
  const float g = param.gravity;
  const float kdd = param.depositionRateDebris;
  const float kds = param.suspensionRateDebris;
  const float tau_y = param.yieldStress;
  const float lenL = len2(scale.x, scale.y);
  const float source_d = st.s0;

  float px = st.px, py = st.py, sx = st.sx, sy = st.sy;
  float att_d = st.att0, att_v = st.att1;
  int ind = st.ind;
  int iter = st.iter & 0x3fffffff;

  bool have = false;  // the window (hw, vl) already holds the current cell (gathered one crossing ahead)
  h5 hw;
  float2 vl;
  hw.c = hw.n0 = hw.p0 = hw.zn = hw.zp = 0.0f; vl = make_float2(0.0f, 0.0f);
  while (true) {
    if (oobf(px, py, nx, ny)) break;
    if (!((unsigned long long)(++iter) < param.maxage)) break;
    const int gx = (int)px, gy = (int)py;
    if (!have) {
      if (!io.relocate(gx)) {
        st.px = px; st.py = py; st.sx = sx; st.sy = sy;
        st.att0 = att_d; st.att1 = att_v;
        st.ind = ind; st.iter = (iter - 1) | (1 << 30);
        export_particle(c.sp, st, gx < c.sp.own_lo ? 0 : 1);
        return;
      }
      const cell_in ci = gather(io, gx, gy, nx, ny, false);
      hw = ci.w; vl = ci.vel;
    }
    const int nind = flattenf(px, py, ny);
    const size_t li = (size_t)nind - io.off;
    if (nind != ind) {
      ind = nind;
      io.deposit_debris((int)li, att_d * source_d, att_v * st.s1, att_v * st.s2, att_d * st.s3, att_d * st.s4,
                        att_d * st.s5);
    }

    const float v_norm = len2(sx, sy);
    const float ux = div_zn(sx, v_norm), uy = div_zn(sy, v_norm);
    const float v_step = stepsize(px, py, ux, uy);
    const float dL = v_step * lenL;
    const float ds = div_zn(dL, v_norm);
    if (v_norm < eps) break;

    const float npx = px + v_step * ux, npy = py + v_step * uy;  // erosion.cu:347, gathered one crossing ahead
    have = false;
    h5 nhw = hw;
    float2 nvl = vl;
    if (!oobf(npx, npy, nx, ny) && (unsigned long long)(iter + 1) < param.maxage) {
      const int ngx = (int)npx, ngy = (int)npy;
      if (io.relocate(ngx)) {
        const cell_in ci = gather(io, ngx, ngy, nx, ny, false);
        nhw = ci.w; nvl = ci.vel;
        have = true;
      }
    }

    const f2 grad = grad_of(hw, scale, param.exitSlope);
    const float debrisHeight = eps + att_d * source_d;
    const float ax = -(g * grad.x) + nu * vl.x;
    const float ay = -(g * grad.y) + nu * vl.y;
    const float decay = nu + tau / debrisHeight;
    const float w = 1.0f / (1.0f + dL * decay);
    sx = w * sx + w * dL * ax;
    sy = w * sy + w * dL * ay;

    const float excessSlope = (len2(grad.x, grad.y) - theta);
    const float excessStress = g * (excessSlope - tau_y / debrisHeight);
    const float shearRate = (excessStress < 0.0f) ? kdd : kds;
    const float decay_d = div_zn(ds * shearRate * excessStress, v_norm);
    const float decay_v = (nu + tau / debrisHeight);
    att_d = att_d * __expf(decay_d);
    att_v = att_v * __expf(-dL * decay_v);
    px = npx;
    py = npy;
    hw = nhw; vl = nvl;
  }
}

// ---- kernels ------------------------------------------------------------------------------

constexpr int kParticleBlock = 64;  // small blocks spread the few, long-lived warps over all SMs (32: +2 %, 128: +7 %)

__global__ void __launch_bounds__(kParticleBlock) k_transport_fluvial(ref_io io, particle_ctx c,
                                                                      const float* waterSource,
                                                                      soilb200_param_t param) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= c.P) return;
  pstate st;
  if (fluvial_init(n, io, c, waterSource, param, st)) fluvial_run(st, io, c, param);
}

__global__ void __launch_bounds__(kParticleBlock) k_transport_debris(ref_io io, particle_ctx c,
                                                                     soilb200_param_t param) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= c.P) return;
  pstate st;
  if (debris_init(n, io, c, param, st)) debris_run(st, io, c, param);
}

// strips: every rank advances every RNG stream (the start positions are the single-GPU ones) but runs only
// the particles that start in its rows.  Left inside the transport kernel that means N-times the blocks, each
// 1/N full - at 8 ranks the grid no longer fits the GPU in one wave and every wave costs a full trajectory.
// So the draw is its own small kernel that appends the rank's starts to a dense list; the transport kernel
// then runs count threads' worth of full warps.  Order within the list is arbitrary (particles are independent).
__global__ void __launch_bounds__(256) k_spawn(curandStateXORWOW_t* rngF, curandStateXORWOW_t* rngD, long long P,
                                               int nx_g, int ny, int own_lo, int own_hi, float2* posF, float2* posD,
                                               int* counts) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = t < 2 * P;
  const int kind = t >= P;
  float px = 0.0f, py = 0.0f;
  bool mine = false;
  if (live) {
    const long long n = t - (kind ? P : 0);
    curandStateXORWOW_t* rng = kind ? rngD : rngF;
    curandStateXORWOW_t rs = rng[n];
    px = 0.5f + curand_uniform(&rs) * float(nx_g - 1);  // erosion.cu:56-59 / :268-271
    py = 0.5f + curand_uniform(&rs) * float(ny - 1);
    rng[n] = rs;
    const int gx = (int)px;
    mine = gx >= own_lo && gx < own_hi;
  }
  // one atomic per warp and kind, not one per particle: P additions to one address take P serial trips to L2
  // (0.3 ms of a 2.2 ms particle phase at P = 65536)
  const unsigned lane = threadIdx.x & 31u;
  for (int k = 0; k < 2; ++k) {  // a warp holds both kinds only where P is not a multiple of 32
    const bool in = mine && kind == k;
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if (!m) continue;
    const int lead = __ffs(m) - 1;
    int base = 0;
    if ((int)lane == lead) base = atomicAdd(counts + k, __popc(m));
    base = __shfl_sync(0xffffffffu, base, lead);
    if (in) (k ? posD : posF)[base + __popc(m & ((1u << lane) - 1u))] = make_float2(px, py);
  }
}

// fused step: both particle systems in ONE launch (they are independent: different flux records,
// different RNG streams), blocks alternating fluvial / debris so that every SM holds warps of both
// kinds and has twice the warps to hide the serial per-particle latency behind
__global__ void __launch_bounds__(kParticleBlock) k_transport_both(fused_io fio, fused_io dio, particle_ctx fc,
                                                                   particle_ctx dc, const float* waterSource,
                                                                   soilb200_param_t param) {
  const bool is_debris = blockIdx.x & 1;
  // a whole map: one particle per thread.  A strip: the grid is sized for the strip's expected share of the starts
  // (launch_transport_both) and a thread takes the list entries n, n + stride, ... - every rank launching all
  // P_total threads' worth of blocks, seven in eight of them empty at eight ranks, cost the particle kernel 0.3 ms
  const int stride = (int)(gridDim.x >> 1) * (int)blockDim.x;   // the launcher keeps P below 2^31
  const int limit = fc.spawn_pos ? __ldg(is_debris ? dc.spawn_count : fc.spawn_count) : (int)fc.P;
  for (int n = (int)(blockIdx.x >> 1) * (int)blockDim.x + (int)threadIdx.x; n < limit; n += stride) {
    pstate st;
    if (!is_debris) {
      if (fluvial_init(n, fio, fc, waterSource, param, st)) fluvial_run(st, fio, fc, param);
    } else {
      if (debris_init(n, dio, dc, param, st)) debris_run(st, dio, dc, param);
    }
  }
}

// particles handed over by a neighbouring strip continue here
__global__ void __launch_bounds__(kParticleBlock) k_transport_resume(fused_io fio, fused_io dio, particle_ctx c,
                                                                     const float* inbox, int n_in,
                                                                     soilb200_param_t param) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_in) return;
  pstate st;
  const float4* src = reinterpret_cast<const float4*>(inbox + 16 * (size_t)n);
  const float4 a = __ldg(src), b = __ldg(src + 1), cc = __ldg(src + 2), d = __ldg(src + 3);
  st.px = a.x; st.py = a.y; st.sx = a.z; st.sy = a.w;
  st.att0 = b.x; st.att1 = b.y; st.att2 = b.z; st.s0 = b.w;
  st.s1 = cc.x; st.s2 = cc.y; st.s3 = cc.z; st.s4 = cc.w;
  st.s5 = d.x; st.s6 = d.y; st.ind = __float_as_int(d.z); st.iter = __float_as_int(d.w);
  if ((st.iter >> 30) == 0) fluvial_run(st, fio, c, param);
  else debris_run(st, dio, c, param);
}

inline strip_t whole_domain(int nx) { return strip_t{0, 0, nx, nx, nullptr, nullptr, nullptr, 0}; }

static strip_t to_strip(const strip_host& h) {
  return strip_t{h.row0, h.own_lo, h.own_hi, h.nx_g, h.outbox_lo, h.outbox_hi, h.counts, h.capacity};
}

// the view a particle of this rank starts with: the rank's own maps
static fused_io own_view(const float* hmap, float* flux, int R, unsigned char* touched, const float* wh,
                         const float* momentum, const float* debrisVel, int rows, int ny, const strip_host& sh,
                         int kind, const void* view) {
  fused_io io;
  io.hmap = hmap; io.R = R; io.touched = touched; io.wh = wh;
  io.rec = flux + (kind ? R : 0);
  io.vel = (const float2*)(kind ? debrisVel : momentum);
  io.off = (size_t)sh.row0 * ny; io.own_lo = sh.own_lo; io.own_hi = sh.own_hi; io.remote = false;
  io.roff = io.off; io.far = false;
  io.my_hmap = io.hmap; io.my_wh = io.wh; io.my_vel = io.vel; io.my_off = io.off;
  io.win_lo = sh.row0; io.win_hi = sh.row0 + rows; io.nx_g = sh.nx_g;
  io.peers = (const peer_dev*)sh.peers; io.n_peers = sh.n_peers; io.self = sh.peer_self; io.kind = kind; io.ny = ny;
  io.nyq = (ny + 1) >> 1;
  io.view = view ? (const float4*)view + (kind ? (size_t)((rows + 1) >> 1) * io.nyq * 4 : 0) : nullptr;
  io.vrow0 = sh.row0;
  io.my_view = io.view; io.my_vrow0 = io.vrow0;
  return io;
}

}  // namespace

// host launchers used by dense.cu's entry points
int launch_transport_fluvial(float* waterFlux, float* massFlux, float* velocityFlux, float* albedoFlux, void* rng,
                             long long P, const float* layers, const float* rainfall, const float* waterHeight,
                             const float* velocity, const float* albedoSource, int nx, int ny, scale3 s,
                             const soilb200_param_t& p, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const ref_io io{(const float2*)layers, waterFlux, massFlux, (float2*)velocityFlux, albedoFlux, waterHeight,
                  (const float2*)velocity, 0};
  const particle_ctx c{(curandStateXORWOW_t*)rng, P, ny, s, albedoSource, whole_domain(nx), nullptr, nullptr};
  k_transport_fluvial<<<blocks_for(P, kParticleBlock), kParticleBlock, 0, st>>>(io, c, rainfall, p);
  count_launch();
  return check_launch();
}

int launch_transport_debris(float* massFlux, float* velocityFlux, float* albedoFlux, void* rng, long long P,
                            const float* layers, const float* velocity, const float* albedoSource, int nx, int ny,
                            scale3 s, const soilb200_param_t& p, cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const ref_io io{(const float2*)layers, nullptr, massFlux, (float2*)velocityFlux, albedoFlux, nullptr,
                  (const float2*)velocity, 0};
  const particle_ctx c{(curandStateXORWOW_t*)rng, P, ny, s, albedoSource, whole_domain(nx), nullptr, nullptr};
  k_transport_debris<<<blocks_for(P, kParticleBlock), kParticleBlock, 0, st>>>(io, c, p);
  count_launch();
  return check_launch();
}

// strips only: draws every particle's start and lists the ones of this rank (k_spawn).  Layout of the scratch:
// int counts[2], pad to 256 B, float2 posF[P], float2 posD[P].  A whole-map model needs none of it.
static bool force_spawn() {  // measurement only: take the k_spawn path on a whole-map strip too
  static const bool on = getenv("SOILB200_DEBUG_FORCE_SPAWN") != nullptr;
  return on;
}

int launch_spawn(void* rngF, void* rngD, long long P, int ny, const strip_host& sh, void* spawn, cudaStream_t st) {
  if (P <= 0 || !spawn || !(sh.own_lo > 0 || sh.own_hi < sh.nx_g || force_spawn())) return SOILB200_OK;
  int* counts = (int*)spawn;
  float2* posF = (float2*)((char*)spawn + 256);
  const cudaError_t e = cudaMemsetAsync(counts, 0, 2 * sizeof(int), st);
  if (e != cudaSuccess) return (int)e;
  k_spawn<<<blocks_for(2 * P, 256), 256, 0, st>>>((curandStateXORWOW_t*)rngF, (curandStateXORWOW_t*)rngD, P, sh.nx_g, ny,
                                                  sh.own_lo, sh.own_hi, posF, posF + P, counts);
  count_launch();
  return check_launch();
}

// flux = per cell {fluvial record [R], debris record [R]}, R = 8 with albedo else 4
int launch_transport_both(const float* hmap, float* flux, int R, unsigned char* touched, void* rngF, void* rngD,
                          long long P, const float* rainfall, const float* waterHeight, const float* velocity,
                          const float* debrisVel, const float* albedoSource, int rows, int ny, scale3 s,
                          const soilb200_param_t& p, const strip_host& sh, void* spawn, const void* view,
                          cudaStream_t st) {
  if (P <= 0) return SOILB200_OK;
  const fused_io fio = own_view(hmap, flux, R, touched, waterHeight, velocity, debrisVel, rows, ny, sh, 0, view);
  const fused_io dio = own_view(hmap, flux, R, touched, waterHeight, velocity, debrisVel, rows, ny, sh, 1, view);
  const strip_t sp = to_strip(sh);
  particle_ctx fc{(curandStateXORWOW_t*)rngF, P, ny, s, albedoSource, sp, nullptr, nullptr};
  particle_ctx dc{(curandStateXORWOW_t*)rngD, P, ny, s, albedoSource, sp, nullptr, nullptr};
  const bool strip = sh.own_lo > 0 || sh.own_hi < sh.nx_g || force_spawn();
  if (spawn && strip) {  // filled by launch_spawn, earlier on this stream
    fc.spawn_pos = (const float2*)((const char*)spawn + 256); fc.spawn_count = (const int*)spawn;
    dc.spawn_pos = fc.spawn_pos + P; dc.spawn_count = fc.spawn_count + 1;
  }
  if (P >= (1ll << 31)) return SOILB200_E_ARG;
  long long threads = P;
  if (fc.spawn_pos) {
    // starts are uniform over the global map: this strip's share is binomial around P * rows / nx_g; a twentieth
    // more and 64 covers it (a larger count only means a second trip for some threads)
    const double share = (double)(sh.own_hi - sh.own_lo) / (double)sh.nx_g;
    threads = (long long)((double)P * share * 1.05) + 64;
    if (threads > P) threads = P;
  }
  k_transport_both<<<2 * blocks_for(threads, kParticleBlock), kParticleBlock, 0, st>>>(fio, dio, fc, dc, rainfall, p);
  count_launch();
  return check_launch();
}

int launch_transport_resume(const float* hmap, float* flux, int R, unsigned char* touched, const float* inbox, int n_in,
                            long long P, const float* waterHeight, const float* velocity, const float* debrisVel,
                            int rows, int ny, scale3 s, const soilb200_param_t& p, const strip_host& sh,
                            const void* view, cudaStream_t st) {
  if (n_in <= 0) return SOILB200_OK;
  const fused_io fio = own_view(hmap, flux, R, touched, waterHeight, velocity, debrisVel, rows, ny, sh, 0, view);
  const fused_io dio = own_view(hmap, flux, R, touched, waterHeight, velocity, debrisVel, rows, ny, sh, 1, view);
  const particle_ctx c{nullptr, P, ny, s, nullptr, to_strip(sh), nullptr, nullptr};
  k_transport_resume<<<blocks_for(n_in, kParticleBlock), kParticleBlock, 0, st>>>(fio, dio, c, inbox, n_in, p);
  count_launch();
  return check_launch();
}

}  // namespace sb
