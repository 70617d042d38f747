/* soilb200.h - C ABI of the B200-native erosion / flow-routing hot path.
 *
 * Drop-in boundary for erosiv/soillib (reference @ 9b9e0871): every entry point
 * below replaces one free function of `namespace soil` that the reference's
 * nanobind lambdas call (python/source/model.cpp:157-421).  Plain pointers and
 * sizes only; all tensor pointers are DEVICE pointers in the reference's own
 * layout (float32 / int32, row-major, `flatten(x,y) = x*ny + y`, trailing
 * vector dimension interleaved: layers (nx,ny,2) = {bedrock, sediment},
 * momentum (nx,ny,2), albedo (nx,ny,3)).  `stream` is a cudaStream_t passed as
 * void* (NULL = default stream).  Calls are asynchronous and never synchronise
 * the device.  No entry point allocates unless its comment says so.
 *
 * Return value: 0 = ok; < 0 = argument error (SOILB200_E_*); > 0 = cudaError_t.
 * There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef SOILB200_H
#define SOILB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SOILB200_OK 0
#define SOILB200_E_ARG (-1)       /* null pointer / non-positive size            */
#define SOILB200_E_EDGE (-2)      /* edge is neither SOILB200_D4 nor SOILB200_D8 (graph.cu:88) */
#define SOILB200_E_WORKSPACE (-3) /* workspace missing or too small              */
#define SOILB200_E_LIMIT (-4)     /* size beyond a documented fixed limit        */

/* soil::edge_t, graph.hpp:11-14.  D8 neighbour order (graph.hpp:35-44):
 * (-1,0),(0,-1),(0,1),(1,0),(-1,-1),(-1,1),(1,-1),(1,1); D4 = first four. */
#define SOILB200_D4 0
#define SOILB200_D8 1

/* soil::param_t, erosion.hpp:17-58 - same field order, same 112-byte layout. */
typedef struct soilb200_param {
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
} soilb200_param_t;

/* Fills *p with the reference defaults (erosion.hpp:20-56). */
void soilb200_param_default(soilb200_param_t* p);

/* Library / device introspection. */
const char* soilb200_version(void);
int soilb200_device_count(void);
/* bytes of one RNG state: silt::rng == curandStateXORWOW (48), SURVEY App. A-4 */
int soilb200_rng_state_bytes(void);
/* number of kernels this library has launched since load (bench bookkeeping) */
uint64_t soilb200_launch_count(void);
/* Per-phase device timing of soilb200_erode_step (bench bookkeeping; no reference counterpart,
 * the reference's soil::timer is wall-clock only, util/timer.hpp:15-69).  While enabled, every
 * step records CUDA events on the caller's stream around its particle kernels and around its
 * dense kernel; _read synchronises the device, returns the summed milliseconds, the number of
 * steps and the number of cells particles touched, and resets the counters. */
int soilb200_profile_enable(int on);
int soilb200_profile_read(double* particle_ms, double* dense_ms, int64_t* steps, uint64_t* hot_segments);

/* Self-test of an arithmetic shortcut: the kernels divide diagonal slopes by sqrtf(2.0f)
 * (graph.cu:56) with a 3-instruction correctly-rounded sequence instead of the generic IEEE
 * division.  This runs all 2^32 binary32 inputs through both and writes the number of
 * differing results (must be 0) to *mismatches_device (a device uint64). */
int soilb200_selftest_div_sqrt2(uint64_t* mismatches_device, void* stream);
/* The same for the divisions of the fused erosion step by its launch constants (cell scale, normalisation length;
 * erosion.cu:160-185, 526-547, erosion_map.cu:131-157): a power-of-two divisor becomes an exact multiplication, any
 * other divisor between 2^-20 and 2^20 a multiplication by its correctly rounded reciprocal plus a residual
 * correction.  Runs all 2^32 numerators through that and through the IEEE division for `divisor`, writes the number
 * of differing results (must be 0) to *mismatches_device and the kind chosen for the divisor (0 plain division,
 * 1 power of two, 2 reciprocal + correction) to *kind (host, may be NULL). */
int soilb200_selftest_div_const(float divisor, uint64_t* mismatches_device, int32_t* kind, void* stream);

/* ---- flow graph (source/soillib/model/graph/graph.cu) ------------------- */

/* soil::direction, graph.cu:201-264: out[n] = neighbour code k of steepest descent, -1 if none. */
int soilb200_direction_f32(const float* height, int32_t* out, int nx, int ny, int edge, void* stream);

/* soil::steepest, graph.cu:27-91: out[n] = flat index of the receiver, -1 if none. */
int soilb200_steepest_f32(const float* height, int32_t* out, int nx, int ny, int edge, void* stream);

/* silt.seed(rng, seed, offset) / __seed, graph.cu:97-101: curand_init(seed, first+i, offset, &state[i]). */
int soilb200_seed_xorwow(void* states, int64_t n, uint64_t seed, uint64_t first_subsequence, uint64_t offset,
                         void* stream);

/* soil::random_weighted, graph.cu:103-195.  The reference allocates and seeds a
 * 48 B/cell curandState tensor on every call.  Here `ws` (optional, may be NULL)
 * is caller-owned scratch of soilb200_random_weighted_ws(nx,ny) bytes that keeps
 * the per-cell generator state between calls: a call with the same seed and
 * `offset` one past the previous call's advances each stream by one draw, which
 * is bit-identical to re-seeding (curand_init(s,n,k) + 1 draw == curand_init(s,n,k+1)).
 * Before its first use the caller zeroes the first 64 bytes of `ws`. */
size_t soilb200_random_weighted_ws(int nx, int ny);
int soilb200_random_weighted_f32(const float* height, int32_t* out, int nx, int ny, int edge, uint64_t seed,
                                 uint64_t offset, float temperature, void* ws, size_t ws_bytes, void* stream);
/* The same on a block of rows cut out of a larger map (row strips): cell n of the block draws from the
 * stream of global cell first_cell + n (curand_init(seed, first_cell + n, offset)), so that the strips of a
 * map make the decisions the undivided map makes.  Receivers are flat indices into the BLOCK.
 * soilb200_random_weighted_f32 is this with first_cell = 0. */
int soilb200_random_weighted_strip_f32(const float* height, int32_t* out, int nx, int ny, int edge, uint64_t seed,
                                       uint64_t offset, float temperature, uint64_t first_cell, void* ws,
                                       size_t ws_bytes, void* stream);

/* soil::accumulate / soil::accumulate_decay, graph.cu:321-593.  decay == NULL
 * selects accumulate (edge weight 1); otherwise per-cell decay with the
 * reference's slot rule (my_decay, graph.cu:382-420).  `ws` is caller-owned
 * scratch: soilb200_accumulate_ws(nx,ny,edge) = 12 bytes per cell without decay,
 * soilb200_accumulate_decay_ws = 20 bytes per cell with it (the reference
 * allocates 2*N*(8+8K) bytes internally).  With that much scratch the sums are
 * resolved in 64x64 shared-memory tiles; with at least 4 bytes per cell by a global
 * dependency walk (same sums, ~8x slower at 8192^2, and the only path for maps
 * narrower than 3 columns); less than 4 bytes per cell is SOILB200_E_WORKSPACE. */
size_t soilb200_accumulate_ws(int nx, int ny, int edge);
size_t soilb200_accumulate_decay_ws(int nx, int ny, int edge);
int soilb200_accumulate_f32(const int32_t* graph, const float* source, const float* decay, float* out, int nx,
                            int ny, int edge, void* ws, size_t ws_bytes, void* stream);

/* For each of n_starts cells (flat indices into an nx x ny graph of receiver indices, -1 = none): the last cell
 * of its flow path, i.e. the first cell on the path that has no receiver (out[i]; -1 for an invalid start).
 * No counterpart function in the reference; it is the piece strip-partitioned accumulation needs on top of
 * soilb200_accumulate_f32 (which exit of the strip does an inflowing boundary cell drain to). */
int soilb200_graph_terminal_i32(const int32_t* graph, const int32_t* starts, int32_t* out, int n_starts, int nx, int ny,
                                void* stream);

/* Accumulation for callers that cannot allocate device memory.  (No counterpart in the
 * reference: soil::accumulate allocates all of its scratch itself, graph.cu:539-550.)
 *
 * Same inputs and outputs as soilb200_accumulate_f32, but without the workspace argument (same
 * sums: bit-identical to that call with decay, with 4 bytes of workspace per cell or with
 * integer-valued sources; the tiled path that call takes with its full workspace and no decay adds
 * the same terms in another order, so real-valued sources then agree to float rounding):
 * the per-cell pending-donor counters live in a fixed device array that is part of the
 * library image, SOILB200_ACCUMULATE_STATIC_MAX_CELLS ints = 16 MiB on every GPU that loads the
 * library, for the life of the process.  Meant for hosts that give the library nothing but
 * pointers to their own maps and forbid cudaMalloc in their processing loop (it locks and
 * synchronises the whole device), or that have no allocator to hand scratch from.  The call itself
 * performs no cudaMalloc / cudaFree and uses no allocating library routine.
 *
 *  - nx*ny > SOILB200_ACCUMULATE_STATIC_MAX_CELLS returns SOILB200_E_LIMIT (use the workspace
 *    variant for larger maps).
 *  - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *    that device current); make that first call before starting any stream capture.
 *  - All calls on a device share the one array.  Its previous contents never matter (every
 *    counter a call reads is written by that call first), but two calls whose kernels overlap on
 *    the GPU - different streams, or different host threads - race on it: serialise them. */
#define SOILB200_ACCUMULATE_STATIC_MAX_CELLS (4 * 1024 * 1024)
int soilb200_accumulate_static_f32(const int32_t* graph, const float* source, const float* decay, float* out,
                                   int nx, int ny, int edge, void* stream);

/* ---- gradient family, slope, blur ------------------------------
 * The reference's versions allocate and return their result (and gaussian_blur clobbers and returns
 * its argument); here the caller provides the output.  scale = {scale.x, scale.y}. */

/* soil::gradient, grad.cu:22-98: out (nx,ny,2) = central difference per axis, falling back to the
 * backward, then the forward one-sided difference, then 0 (neighbours outside the map are NaN). */
int soilb200_gradient_f32(const float* height, float* out, int nx, int ny, const float scale[2], void* stream);

/* soil::negslope, grad.cu:102-140: out (nx,ny) = length of the per-axis steepest downhill slope. */
int soilb200_negslope_f32(const float* height, float* out, int nx, int ny, const float scale[2], void* stream);

/* soil::laplacian, grad.cu:137-203: 9-point Laplacian with boundary continuation of a map with
 * `channels` = 1 or 2 interleaved components (other values: SOILB200_E_ARG; the reference silently
 * returns an uninitialised tensor).  in != out. */
int soilb200_laplacian_f32(const float* in, float* out, int nx, int ny, int channels, const float scale[2],
                           void* stream);

/* soil::slope, graph.cu:270-311: (h[next] - h[n]) / |scale * (pos(next) - pos(n))| along the flow
 * graph `flow` (receiver indices as written by soilb200_steepest_f32), 0 where there is none. */
int soilb200_slope_f32(const float* height, const int32_t* flow, float* out, int nx, int ny, const float scale[2],
                       void* stream);

/* soil::gaussian_blur, filter.cu:24-91: separable 33-tap Gaussian (window +-16, clamped indices,
 * weights not renormalised at the edge), first along dim 0 (inout -> tmp), then along dim 1
 * (tmp -> inout): like the reference's, the argument is overwritten and holds the result.
 * `tmp`: caller scratch of the same size (the tensor the reference allocates).  channels 1 or 2. */
int soilb200_gaussian_blur_f32(float* inout, float* tmp, int nx, int ny, int channels, float sigma, void* stream);

/* The multi-flow loop of example/dem_multiflow.py:43-49 as one call (no counterpart function in the
 * reference: the script composes it, copying every sample to the host).  For k in
 * [k_first, k_first + k_count): graph_k = random_weighted(height, edge, seed, offset = k, T),
 * acc_k = accumulate(graph_k, source), out += (double)(acc_k / (float)k_total) - the script's
 * arithmetic (binary32 quotient, binary64 sum, k ascending), so one call over all K samples
 * reproduces its mean bit for bit.  add_to_out == 0 overwrites `out` (nx*ny doubles), != 0 adds to
 * what it holds.  Replicas on several GPUs take disjoint k ranges and sum their `out` maps.
 * `ws`: soilb200_multiflow_ws(nx,ny,edge) bytes, 256-byte aligned, contents irrelevant (graph, one
 * accumulation map, the per-cell generator streams, the accumulate scratch: 44 B/cell).
 * Everything is enqueued on `stream`; nothing synchronises. */
size_t soilb200_multiflow_ws(int nx, int ny, int edge);
int soilb200_multiflow_f64(const float* height, const float* source, double* out, int nx, int ny, int edge,
                           uint64_t seed, uint64_t k_first, uint64_t k_count, uint64_t k_total, float temperature,
                           int add_to_out, void* ws, size_t ws_bytes, void* stream);

/* ---- erosion (source/soillib/model/path/erosion.cu) --------------------- */

/* soil::transport_fluvial, erosion.cu:189-239; argument order of erosion.hpp:69-84.
 * Particles accumulate into the *_track tensors (never zeroed here, App. B-9),
 * then the normalise pass overwrites discharge, mass, momentum and turns
 * albedo_transport into a colour.  n_particles = rng.elem(). */
int soilb200_transport_fluvial(const float* layers, const float* rainfall, float* discharge, float* discharge_track,
                               float* mass, float* mass_track, float* momentum, float* momentum_track,
                               const float* albedo_bedrock, float* albedo_transport, const float* albedo_surface,
                               void* rng, int64_t n_particles, int nx, int ny, const float scale[3],
                               const soilb200_param_t* param, void* stream);

/* soil::transport_debris, erosion.cu:395-436; argument order of erosion.hpp:86-98. */
int soilb200_transport_debris(const float* layers, float* velocity, float* velocity_track, float* mass,
                              float* mass_track, const float* albedo_bedrock, float* albedo_transport,
                              const float* albedo_surface, void* rng, int64_t n_particles, int nx, int ny,
                              const float scale[3], const soilb200_param_t* param, void* stream);

/* soil::mass_transfer, erosion.cu:576-611; argument order of erosion.hpp:104-119.
 * The four albedo pointers may be NULL together (albedo mixing skipped). */
int soilb200_mass_transfer(float* delta, const float* layers, const float* uplift, const float* discharge,
                           const float* mass, const float* momentum, const float* debris,
                           const float* momentum_debris, const float* albedo_bedrock,
                           const float* albedo_transport_fluvial, const float* albedo_transport_debris,
                           float* albedo_surface, int nx, int ny, const float scale[3],
                           const soilb200_param_t* param, void* stream);

/* soil::mass_creep, erosion.cu:712-727. */
int soilb200_mass_creep(float* delta, const float* layers, int nx, int ny, const float scale[3],
                        const soilb200_param_t* param, void* stream);

/* soil::layer_merge, erosion.cu:747-757. */
int soilb200_layer_merge(float* height, const float* layers, int nx, int ny, void* stream);

/* soil::solve_uniform, path.cu:169-217: uniform-sampling Monte-Carlo estimator of a
 * linear conservation law.  flow (nx,ny,2) [m/s], source (nx,ny,components) [X/s], decay (nx,ny) [1/s];
 * `rng`: n_samples XORWOW states (one sample per state, each advanced by two draws); `flux`
 * (nx,ny,components) is zeroed, filled by the samples and normalised in place - the tensor the reference
 * allocates and returns.  components = 1 or 2 (others: SOILB200_E_ARG; the reference returns the zeroed
 * tensor).  epsilon 1e-16 and the nx + ny step bound are the reference's (path.cu:193-194).  Unlike the
 * reference (path.cu:212) nothing synchronises. */
int soilb200_solve_uniform_f32(const float* flow, const float* source, const float* decay, void* rng,
                               int64_t n_samples, float* flux, int nx, int ny, int components, const float scale[2],
                               int64_t count, void* stream);

/* ---- albedo generators, erosion.cu:759-919 ---------------------
 * Colour maps are (nx, ny, 3) interleaved, `layers` is (nx, ny, 2). */

/* soil::albedo_layer, erosion.cu:878-898: albedo = blend * bedrock + (1 - blend) * min(1, sediment + shift),
 * blend = 1 / (1 + scale_sediment * sediment thickness). */
int soilb200_albedo_layer(float* albedo, const float* albedo_bedrock, const float* albedo_sediment, const float* layers,
                          int nx, int ny, float scale_sediment, const float shift_sediment[3], void* stream);

/* soil::albedo_stratum, erosion.cu:828-855: bedrock colour = color_a / color_b alternating with
 * floor(max(age * param.uplift * uplift - bedrock * scale.z, 0) / freq). */
int soilb200_albedo_stratum(float* albedo_bedrock, const float* uplift, const float* layers, int nx, int ny,
                            const float scale[3], const soilb200_param_t* param, const float color_a[3],
                            const float color_b[3], float age, float freq, void* stream);

/* soil::albedo_discharge, erosion.cu:900-917: in-place extinction blend towards color_discharge,
 * blend = scale * (1 - __expf(-extinction * max(0, discharge))). */
int soilb200_albedo_discharge(float* albedo, const float* discharge, int nx, int ny, const float color_discharge[3],
                              float extinction, float scale, void* stream);

/* soil::op::normal, op/normal.hpp:18-39: out (nx,ny,3) = normalize(-gx,-gy,1).
 * The gradient stencil body lives in the missing silt dependency; it follows the
 * specification in SURVEY.md App. A-8 (parity unpinned). */
int soilb200_normal_f32(const float* height, float* out, int nx, int ny, const float scale[3], void* stream);

/* soil::noise, op/noise.hpp:42-56 (FastNoiseLite 1.0.1 OpenSimplex2 fBm).  The
 * reference fills a HOST tensor in a serial loop; `out` here may be a host
 * pointer (on_device = 0, computed on the calling thread) or a device pointer
 * (on_device = 1, one kernel).  Both produce the same bits. */
int soilb200_noise_f32(float* out, int nx, int ny, float frequency, int octaves, float gain, float lacunarity,
                       float seed, float ext_x, float ext_y, int on_device, void* stream);

/* One whole erosion step, replacing what `soil.erode(model, data, track, param, 1)`
 * did in the previous soillib API (example/erosion_gpu.py:103-106; at HEAD the
 * step is a composition of the ops above): zero fluxes,
 * fluvial + debris particle transport, normalise, mass transfer, optional creep,
 * apply, merge - as ONE particle kernel (both particle systems) and ONE fused
 * dense pass.
 *
 * State maps keep the reference layout.  `layers` and `layers_next` ping-pong: the
 * step reads `layers`, writes `layers_next`; the caller swaps them.  `height` must
 * hold bedrock + sediment of `layers` on entry (soilb200_layer_merge) and holds that
 * of `layers_next` on exit; the particles read it.  albedo_bedrock / albedo_surface
 * may be NULL together.  lrate != 1 applies x <- (1-lrate) x + lrate x_new to the five
 * transported maps (the reference's normalise kernels overwrite, i.e. lrate == 1).
 *
 * The flux accumulators (the reference's *_track tensors) are private scratch in a
 * packed layout: `flux` is soilb200_erode_flux_bytes(nx,ny,albedo) bytes, 16-byte
 * aligned, all zero before the first step; `touched` is soilb200_erode_ws(nx,ny)
 * bytes of tags (one bit per cell), zero before the first step.
 *   - flux_prev == NULL (in-place mode): the dense pass reads and re-zeroes `flux`
 *     and clears `touched`.  Simple, but read-then-zero of the same lines runs ~5x
 *     below copy bandwidth on B200.
 *   - flux_prev != NULL (ping-pong mode): this step's particles fill `flux` (clean
 *     on entry), the dense pass only READS it, re-zeroes - with pure stores - the
 *     cells of `flux_prev` that `touched_prev` (the previous step's tags) marks, and
 *     clears `touched_prev`.  After the call the caller swaps {flux, touched} with
 *     {flux_prev, touched_prev}.  Both buffers start all zero.
 * `mass_sum` (optional, 2 doubles) receives sum(bedrock), sum(sediment) of layers_next. */
typedef struct soilb200_model {
  int32_t nx, ny;
  float scale[3];
  int32_t with_creep;
  float *layers, *layers_next, *height;
  const float *rainfall, *uplift;
  float *discharge, *mass, *momentum, *debris, *debris_momentum;
  const float* albedo_bedrock;
  float* albedo_surface;
  void *rng_fluvial, *rng_debris;
  int64_t n_particles;
  void *flux, *touched;
  void *flux_prev, *touched_prev;
  double* mass_sum;
  /* ---- row strips (multi-GPU; all zero = the arrays are the whole map) ----
   * The map is cut into contiguous row strips, one per GPU.  The arrays of this model then hold
   * global rows [global_row0, global_row0 + nx) of a global_nx-row map: the rows this rank owns,
   * [own_row_lo, own_row_hi) in GLOBAL numbering, plus one halo row on each side that has a
   * neighbour (valid in `layers` and `height` on entry; the caller refreshes them from the
   * neighbours after every step).  The dense pass updates owned rows only and treats only the
   * global first/last row as the map edge.  Every rank draws the start of all n_particles
   * particles from identical rng states and runs those that start in its rows; a particle that
   * moves into a neighbour's rows is appended (64 bytes, layout private) to outbox_lo (towards
   * lower rows) or outbox_hi, outbox_counts[0/1] counting them (zeroed by the caller; a count
   * above outbox_capacity means records were dropped).  The caller ships the records to the
   * neighbour, which continues them with soilb200_erode_resume; repeat until no rank has any.
   * `phases` selects what one call does (0 = both): particles are launched first, the dense pass
   * after all hand-offs are done. */
  int32_t global_nx, global_row0, own_row_lo, own_row_hi;
  void *outbox_lo, *outbox_hi;
  int32_t* outbox_counts;
  int32_t outbox_capacity;
  int32_t phases;
  /* Peer-mapped strips (one NVSwitch box): `peers` is a DEVICE array of n_peers soilb200_peer_t, entry r
   * describing rank r's rows and the device addresses - valid on THIS GPU, e.g. opened from CUDA IPC handles -
   * of its height, discharge, momentum, debris_momentum, flux and tag arrays; peer_self is this rank's entry.
   * With it a particle that moves into another rank's rows simply keeps going: it gathers and deposits
   * through that rank's pointers over NVLink (deposits as 4-byte reductions), and no outbox is used.  The
   * caller must then order the phases across ranks: every rank's particle kernel before any rank's dense
   * pass, every dense pass (and halo refresh) before the next step's particles. */
  const void* peers;
  int32_t n_peers, peer_self;
  /* Optional scratch of soilb200_erode_pre_bytes(nx,ny) bytes (16-byte aligned; contents need not be
   * initialised).  With it the gradient and creep term of the pre-update layers are computed by a separate
   * kernel on a library-owned side stream WHILE the particle kernel runs (which leaves ~80 % of the issue slots
   * idle), and the dense pass reads them back instead of recomputing: same results, shorter critical path.
   * The first such call on a device creates the side stream and two events. */
  void* pre;
  /* Optional scratch of soilb200_erode_spawn_bytes(n_particles) bytes (256-byte aligned; contents need not be
   * initialised), used only when the model is a strip of a larger map: the start positions of all particles
   * are then drawn by a small kernel that lists the ones starting in this rank's rows, and the particle kernel
   * runs that dense list - instead of n_particles threads per rank of which (N-1)/N exit at once, which at 8
   * ranks no longer fits the GPU in one wave of blocks.  Same particles, same trajectories. */
  void* spawn;
  /* Optional: the dense pass of a strip in several calls (all zero = one call over all owned rows).  A call with
   * SOILB200_PHASE_DENSE and dense_row_lo < dense_row_hi updates only the owned rows inside that GLOBAL row range.
   * A step's dense calls must cover every owned row exactly once; every call but the last adds
   * SOILB200_PHASE_DENSE_MORE (the spent tags are cleared by the last call only), every call but the first adds
   * SOILB200_PHASE_DENSE_CONT (mass_sum keeps accumulating, the precomputed gradient/creep record is already joined).
   * Use: a strip runs the rows its neighbours keep as halo first and ships them while the interior is computed. */
  int32_t dense_row_lo, dense_row_hi;
  /* Optional scratch of soilb200_erode_view_bytes(nx,ny) bytes (16-byte aligned): the packed particle view.  What a
   * particle gathers at a crossing (height of the cell and its four neighbours, velocity, water height) then comes
   * from one 16-byte record per cell and particle system - fluvial {height, discharge, momentum.x, momentum.y}, debris
   * {height, debris_momentum.x, debris_momentum.y, debris} - stored in 2 x 2-cell quads (one 64-byte DRAM request), which
   * the dense pass writes INSTEAD of the maps height, discharge, momentum, debris and debris_momentum: with `view` set
   * the step neither reads nor updates those five maps.  soilb200_erode_import fills the view from them (before the
   * first step, and after the caller edited them), soilb200_erode_export writes them from the view (when the caller
   * wants to read them).  Needs flux_prev / touched_prev (ping-pong); otherwise SOILB200_E_ARG.  Strips: every rank's view covers its allocated rows (halo rows included), the halo rows are
   * refreshed by the caller like the maps they replace (whole quad rows: global_row0 and the halo depth even). */
  void* view;
} soilb200_model_t;

/* Lets kernels on the current device dereference memory of `peer_device` (cudaDeviceEnablePeerAccess;
 * already-enabled is not an error; SOILB200_E_LIMIT if the two GPUs have no peer path).  Needed once per
 * peer before `peers` pointers that live on other GPUs are used. */
int soilb200_enable_peer_access(int peer_device);

/* CUDA IPC plumbing for the peer table (one process per GPU).  _export: the 64-byte handle of the
 * allocation that contains `ptr` and ptr's offset in it (callers using a sub-allocating allocator get the
 * right base).  _open: maps such an allocation into this process for the CURRENT device, enabling peer
 * access to the exporting GPU; an allocation may be opened once per process - cache the base. */
int soilb200_ipc_export(const void* ptr, unsigned char handle[64], uint64_t* offset);
int soilb200_ipc_open(const unsigned char handle[64], void** base);
/* Unmaps an allocation opened with soilb200_ipc_open (cudaIpcCloseMemHandle); call it before the exporter frees it. */
int soilb200_ipc_close(void* base);

typedef struct soilb200_peer {
  int32_t own_row_lo, own_row_hi, global_row0, rows;
  const float *height, *discharge, *momentum, *debris_momentum;
  float* flux;
  void* touched;
  const void* view; /* the rank's packed particle view (NULL unless every rank runs with one) */
  const void* reserved;
} soilb200_peer_t;

#define SOILB200_PHASE_PARTICLES 1
#define SOILB200_PHASE_DENSE 2
#define SOILB200_PHASE_DENSE_MORE 4 /* further dense calls of this step follow */
#define SOILB200_PHASE_DENSE_CONT 8 /* an earlier dense call of this step has run */

size_t soilb200_erode_ws(int nx, int ny);
size_t soilb200_erode_flux_bytes(int nx, int ny, int albedo);
size_t soilb200_erode_pre_bytes(int nx, int ny);
size_t soilb200_erode_spawn_bytes(int64_t n_particles);
size_t soilb200_erode_view_bytes(int nx, int ny);
/* maps -> view (all allocated rows; heights are bedrock + sediment of `layers`) / view -> maps.  No reference
 * counterpart: the view is this library's private layout of maps the reference keeps row-major. */
int soilb200_erode_import(const soilb200_model_t* model, void* stream);
int soilb200_erode_export(const soilb200_model_t* model, void* stream);
int soilb200_erode_step(const soilb200_model_t* model, const soilb200_param_t* param, void* stream);
/* Continues n_in particles received from a neighbouring strip (records as written to an outbox;
 * 16-byte aligned).  They deposit into `flux`, may leave again through the outboxes.  No reference
 * counterpart: soillib is single-GPU. */
int soilb200_erode_resume(const soilb200_model_t* model, const soilb200_param_t* param, const void* inbox,
                          int n_in, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SOILB200_H */
