// kernels.cuh — parameter blocks and launchers of the sm_100a kernels.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "device_math.cuh"

namespace doapp {

constexpr int kMaxJoints = 5;
constexpr uint64_t kKeyNone = 0xFFFFFFFFFFFFFFFFull;
// cost_kernel_pk with 2 or 4 parts per sample: 32 KB of per-waypoint sums per CTA at most
constexpr uint32_t kPkMaxPartN = 256;

enum NoiseMode : int { NOISE_ZERO = 0, NOISE_EXTERNAL = 1, NOISE_PHILOX = 2 };
// PEND_SEED: the previous launch scored the unperturbed trajectories, its key seeds cur_hi
// (generate_trajectories.cu:180-191); PEND_KEEP_BEST: the winner replaces the trajectory if better
// PEND_WEIGHTED: the weighted update vector of the previous iteration (update_vector_kernel) is
// subtracted from the trajectory (generate_trajectories.cu:117-142,242-251)
enum PendMode : int { PEND_NONE = 0, PEND_KEEP_BEST = 1, PEND_SEED = 2, PEND_WEIGHTED = 3 };
// fixed-point scale of the order-independent update-vector sums (2^28) and the largest term
constexpr float kUpdateScale = 268435456.0f;
constexpr float kUpdateTermMax = 4096.0f;

// Device view of the voxel grid (dense, index ix + nx*iy + nx*ny*iz).
struct GridDev {
  const void *cells;
  int kind;  // doapp_grid_kind
  int frame; // doapp_grid_frame
  uint32_t nx, ny, nz;
  // CDG frame (cpu_distance_grid.cu:69-82), all fp32
  float res, x_off, y_off, z_max;
  // OG frame (occupancy_grid.cpp:29-40), fp64
  double res_d, half_d;
  // 1: the reciprocal-multiply index quotient was verified equal to the IEEE one for
  // every dividend this frame can produce (doapp_capi.cu: verify_fast_index)
  int fast_idx;
  float res_rcp;
  // fast_idx only: a point is inside iff, per axis, the BIT PATTERN of x + x_off (y + y_off,
  // z + 0) is below these thresholds as an unsigned integer: negative sums and NaN have the
  // top bits set, and the threshold is the smallest dividend whose voxel index reaches n, so
  // one compare replaces the two float bounds tests and the index test (doapp_grid_alloc
  // establishes and checks the equivalence)
  uint32_t tx_bits, ty_bits, tz_bits;
};

// URDF-correct chain with per-link collision spheres (doapp_fk_model in
// doapp.h).  Spheres are sorted by link; first[l] .. first[l+1] are link l's; order[q] is the
// caller's index of sorted sphere q; thr = radius + collision_distance (fp32 add, host).
constexpr int kFkMaxSpheres = 16;
struct FkModelDev {
  float origin[5][3];
  int axis[5];
  uint32_t first[6];
  uint32_t order[kFkMaxSpheres];
  float centre[kFkMaxSpheres][3];
  float thr[kFkMaxSpheres];
};

struct CostParams {
  // fk_on = 0: the chain of generate_trajectories.cu:397-463.  1: the model below, carried by
  // value so the kernels read it from the constant bank (warp-uniform operands, no loads)
  int fk_on;
  FkModelDev fk;
  // trajectories, structure-of-arrays: traj[c][j][npad]
  const float *traj;
  const float *start; // [d] device
  const float *noise; // external buffer or nullptr
  uint64_t noise_m;   // samples per climber in the external buffer
  uint64_t noise_t0;  // global sample index of the buffer's first column
  int noise_mode;
  uint32_t k, m, n, npad, n_noise;
  uint32_t iteration, seed_lo, seed_hi;
  float sigma;
  uint64_t sample_offset, m_global;
  float dt, smooth_scale, coll_dist, limit_weight;
  float lim_min[kMaxJoints], lim_max[kMaxJoints];
  int use_limits;
  int collision_mode, select_mode;
  float rpy_s, rpy_c; // deterministic sin/cos of the fixed -pi/2 pitch (gt.cu:456)
  GridDev grid;
  float *scores;                 // [k*m] or nullptr
  int32_t *collisions;           // [k*m] or nullptr
  unsigned long long *keys;      // [k] packed argmin keys for the update rule, or nullptr
  unsigned long long *keys_float; // [k] packed (float bits, t) keys for the host-style argmin
  dm::FastDiv dt_div, res_div; // divisor + RN(1/divisor)
  int pk_mode; // packed two-samples-per-thread kernel: -1 auto, 0 never, 1 whenever legal (DOAPP_PK)
  int fast_ok; // 1: the reciprocal dt divide was verified and dt is in [kFastDtLo, kFastDtHi]
  // pending keep-best of the previous launch, applied by this one while it stages the
  // trajectory (see cost_kernel): winner key per climber, the noise that launch drew, the
  // climber's current rounded score in / out, and the other trajectory buffer to write
  int pend_mode;                        // PendMode
  const unsigned long long *pend_keys;  // [k]
  uint32_t pend_iteration;
  int pend_noise_mode;
  const float *pend_noise;              // external buffer of that launch (same layout as noise)
  const uint32_t *cur_hi_in;            // [k]
  uint32_t *cur_hi_out;                 // [k]
  uint8_t *found_better;                // [k]
  float *traj_next;                     // [k][d][npad] or nullptr
  unsigned long long *keys_reset;       // [k] key buffer to re-arm for a later launch, or nullptr
  // PEND_WEIGHTED: integer sums [k][8] left by update_vector_kernel (slot 7 = poisoned terms),
  // the gain of generate_trajectories.cu:140, and the sum buffer to zero for the next one
  const long long *pend_usum;
  long long *usum_reset;
  float update_gain;
  // Multi-GPU argmin merge without a collective launch (xr_nranks > 1): the pending keys of the
  // previous iteration are not read from pend_keys but from this rank's peer mailbox, where every
  // rank's xchg_send_kernel stored its key over NVLink: xr_recv[(c * 8) + r] is rank r's key for
  // climber c (kKeyNone until it arrives); the minimum is the merged key.
  uint32_t xr_nranks;
  const unsigned long long *xr_recv;
  unsigned int *xr_error;          // set when a peer's key did not arrive within xr_timeout_ns
  unsigned long long xr_timeout_ns;
  // launch geometry (filled by launch_cost): 2^parts_log2 threads per sample, each owning
  // wpp consecutive waypoints; 2^pw_log2 of them per CTA, the rest across a cluster
  uint32_t parts_log2, pw_log2, wpp;
};

// Persistent planner loop (plan_persistent_kernel): the whole keep-best loop of
// generate_trajectories.cu:197-253 in ONE cooperative launch, one grid-wide barrier per
// iteration.  `cost` carries the problem (traj = input half, traj_next = output half,
// pend_keys = the seed pass's keys, cur_hi_out / found_better = final state).
// plan_persistent_kernel: 8 compute warps + 1 helper warp (the grid-wide key exchange)
constexpr int kPersistComputeThreads = 256, kPersistThreads = kPersistComputeThreads + 32;
struct PersistParams {
  CostParams cost;
  uint32_t iters;
  uint32_t t_log2;             // threads per sample = 2^t_log2, each owning cost.wpp waypoints
  unsigned long long *keys3;   // [3][k] rotating per-iteration barrier words {argmin key, arrivals - 1}, 16 B each (all ones at launch)
  const float *noise_base;     // external noise of iteration 0 (or nullptr: Philox)
  uint64_t noise_iter_stride;  // elements between iterations in the external buffer
  // Multi-GPU (nranks > 1): the ranks' argmin keys are merged through peer memory over NVLink
  // inside the kernel, no NCCL call per iteration.  mailbox[r] is rank r's mailbox (own one
  // included), laid out [3][k][nranks]; rank s stores its key of iteration it for climber c in
  // mailbox[r][(it % 3)][c][s] (stride 8 ranks) of every rank r and polls its own.  result [3][k] republishes the
  // merged key to the rank's other CTAs.  All slots are kKeyNone when the kernel starts.
  uint32_t nranks, rank;
  unsigned long long *mailbox[8];
  unsigned long long *mailbox_next; // own mailbox set of the NEXT launch, re-armed at kernel start
  unsigned long long *result;
};

// Second-generation persistent planner (plan_pipeline_kernel, kernels_persist2.cu): one CTA per
// SM, warp-specialised (compute / fold / collect / noise), speculation `depth` iterations deep,
// tagged 8-byte key words instead of barriers.
constexpr int kP2Ring = 32;        // iterations in the global entry / mailbox rings (>= 2 * depth + 1)
constexpr int kP2Q = 16;           // iterations in the shared-memory rings (>= depth + 1)
constexpr int kP2MaxDepth = 15;
constexpr int kP2FoldWarps = 2;    // iterations alternate between them
constexpr int kP2NoiseWarps = 3;
constexpr int kP2HelperWarps = kP2FoldWarps + 2 + kP2NoiseWarps; // fold, collect, send, noise
constexpr int kP2MaxComputeWarps = 14;
constexpr int kP2MaxThreads = 32 * (kP2MaxComputeWarps + kP2HelperWarps);
struct Persist2Params {
  CostParams cost;             // traj = input, traj_next = output, pend_keys = the seed pass's keys
  uint32_t iters;
  uint32_t t_log2;             // lanes per sample = 2^t_log2 (<= 64)
  uint32_t wpt;                // waypoints per lane (lane r: waypoints r, r + T, ...)
  uint32_t sb;                 // samples per CTA
  uint32_t depth;              // iterations costed ahead of the last decided one
  unsigned long long *entries; // [kP2Ring][k][gridDim.x] tagged CTA keys (any content at launch
                               // except a valid tag of the first ring turn: memset 0xFF)
  const float *noise_base;     // external noise of iteration 0 (or nullptr: Philox)
  uint64_t noise_iter_stride;
  uint32_t nranks, rank;
  unsigned long long *mailbox[8];   // rank r's mailbox [kP2Ring][k][8] (own one included)
  unsigned long long *mailbox_next; // own mailbox set of the NEXT launch, re-armed at kernel start
  unsigned int *error;              // device word: set when the key exchange timed out
  unsigned long long *stats;        // [8] accepted updates, cost passes of CTA 0, SM clocks by phase (or nullptr)
  unsigned long long timeout_ns;
  uint32_t poll_ns;                 // back-off between polls of the collect / send warps
};
cudaError_t pipeline_plan_config(const CostParams &p, int d, int sm_count, Persist2Params *q, bool *ok);
cudaError_t launch_pipeline_plan(const Persist2Params &q, int d, int sm_count, cudaStream_t st);
uint32_t pipeline_grid_x(const Persist2Params &q);
const void *pipeline_kernel_dim4(int gm, bool lim, bool fkm);
const void *pipeline_kernel_dim5(int gm, bool lim, bool fkm);

struct FinalParams {
  const float *traj; // [k][d][npad]
  const float *start;
  const float *scores;         // [k] float score of every climber's kept trajectory
  const uint8_t *found_better; // [k]
  uint32_t k, n, npad, d;
  float dt;
  // out: [0]=best score, [1..2]=best climber (int64 bits), [4 ..] traj [n][d], then vel [n][d]
  float *out;
};

// compute_update_vector (generate_trajectories.cu:117-133) for every climber of this shard
struct UpdateParams {
  const float *scores;           // [k][m] of the cost launch just run
  const unsigned long long *keys; // [k] CEIL_INT keys of that launch (hi = the block's best_score)
  const float *noise;            // external [k][noise_m][1][d] or nullptr (Philox)
  uint64_t noise_m, noise_t0;
  int noise_mode;
  uint32_t k, m, d, iteration, seed_lo, seed_hi;
  float sigma;
  uint64_t sample_offset;
  long long *usum; // [k][8] accumulated with integer atomics
};

// Depth images -> voxel occupancy in one pass (mapping_node.cpp:89-137 deprojection, :195
// transform to the arm frame, :262-267 crop box, then OccupancyGrid::set per point).
struct DepthCam {
  float cx, cy, kx, ky; // principal point (float, :97-98) and unit_scaling / fx, / fy (:101-103)
  float r[9], t[3];     // camera -> arm_bundle, row-major rotation + translation
};
struct DepthParams {
  const uint16_t *depth; // [cams][h][w], millimetres, 0 = invalid (depth_traits)
  uint32_t cams, w, h;
  DepthCam cam[4];
  float crop_lo[3], crop_hi[3]; // inclusive box in the arm frame (:262-263)
  GridDev grid;
  uint8_t *cells;
  int8_t val;
  unsigned long long *kept; // number of points that passed the crop (may be nullptr)
};

struct ScatterParams {
  const float *xyz;
  uint64_t num_points;
  GridDev grid;
  uint8_t *cells;
  int8_t val;
};

cudaError_t launch_cost(const CostParams &p, int d, int sm_count, cudaStream_t st);
// The cost / planner kernels are compiled once per joint count (kernels_cost.cu, kernels_pk.cu,
// kernels_persist.cu with -DDOAPP_D=4 and -DDOAPP_D=5); launch_cost and the persistent planner
// (kernels_aux.cu) dispatch to these.
#define DOAPP_DIM_CAT2(a, b) a##b
#define DOAPP_DIM_CAT(a, b) DOAPP_DIM_CAT2(a, b)
#define DOAPP_DIM_FN(name) DOAPP_DIM_CAT(name, DOAPP_D)
cudaError_t launch_cost_dim4(const CostParams &p, dim3 grid, unsigned threads, size_t smem,
                             unsigned cluster, cudaStream_t st);
cudaError_t launch_cost_dim5(const CostParams &p, dim3 grid, unsigned threads, size_t smem,
                             unsigned cluster, cudaStream_t st);
cudaError_t launch_cost_pk_dim4(const CostParams &p, int gm, cudaStream_t st);
cudaError_t launch_cost_pk_dim5(const CostParams &p, int gm, cudaStream_t st);
const void *persistent_kernel_dim4(int gm, bool lim, bool fkm);
const void *persistent_kernel_dim5(int gm, bool lim, bool fkm);
cudaError_t launch_verify_div(dm::FastDiv d, uint32_t lo_bits, uint32_t hi_bits, int mode,
                              unsigned long long *bad, int sm_count, cudaStream_t st);
// geometry + co-residency check; *ok = false when the problem does not fit one cooperative wave
cudaError_t persistent_plan_config(const CostParams &p, int d, int sm_count, PersistParams *q,
                                   bool *ok);
cudaError_t launch_persistent_plan(const PersistParams &q, int d, cudaStream_t st);
// Sends this rank's per-climber keys of one iteration into every rank's mailbox slot (one 8-byte
// store per peer and climber over NVLink) after re-arming the own slot of the next iteration
// and, on the first iteration of a plan, the own mailbox set of the next plan.
struct XchgSendParams {
  const unsigned long long *keys;   // [k] this rank's keys of the iteration
  unsigned long long *peer_slot[8]; // rank r's mailbox slot of this iteration: [k][8]
  unsigned long long *rearm_slot;   // own slot of the next iteration [k][8]
  unsigned long long *rearm_set;    // own set of the next plan [3][k][8], or nullptr
  uint32_t k, nranks, rank;
};
cudaError_t launch_xchg_send(const XchgSendParams &p, cudaStream_t st);
cudaError_t launch_update_vector(const UpdateParams &p, int sm_count, cudaStream_t st);
cudaError_t launch_final(const FinalParams &p, cudaStream_t st);
// Occupancy (u8) -> exact Euclidean distance field (f32): three launches (rows, then a
// lower-envelope scan along y and along z); tmp_a / tmp_b hold squared distances (int32).
constexpr int kEdtInf = 1 << 29;
constexpr uint32_t kEdtMaxDim = 768; // scan state of a line lives in shared memory
cudaError_t launch_distance_field(const uint8_t *occ, uint32_t nx, uint32_t ny, uint32_t nz,
                                  float res, int32_t *tmp_a, int32_t *tmp_b, float *out,
                                  cudaStream_t st);
cudaError_t launch_depth_scatter(const DepthParams &p, int sm_count, cudaStream_t st);
cudaError_t launch_scatter(const ScatterParams &p, int sm_count, cudaStream_t st);
cudaError_t launch_gather_probe(const void *cells, uint64_t num_cells, uint32_t cell_bytes,
                                uint32_t blocks, uint32_t per_thread, uint32_t *sink,
                                cudaStream_t st);
cudaError_t launch_lookup(const GridDev &g, const float *xyz, uint64_t count, float coll_dist,
                          long long *cell_index, uint8_t *collides, cudaStream_t st);
cudaError_t launch_noise_fill(const CostParams &p, int d, float *out, cudaStream_t st);
cudaError_t launch_aos_to_soa(const float *aos, float *soa, uint32_t k, uint32_t n, uint32_t npad,
                              uint32_t d, int broadcast_first, cudaStream_t st);
cudaError_t launch_debug_sincos(const float *x, uint64_t count, float *s, float *c,
                                cudaStream_t st);
cudaError_t launch_debug_fk_model(const FkModelDev *model, const float *wp, uint64_t count, int d,
                                  uint32_t num_spheres, float *out, cudaStream_t st);
cudaError_t launch_debug_fk(const float *wp, uint64_t count, int d, float rpy_s, float rpy_c,
                            float *out, cudaStream_t st);
cudaError_t launch_reduce_keys_float(const unsigned long long *keys, uint32_t k,
                                     unsigned long long *out, cudaStream_t st);

// host-side evaluation of the deterministic sin/cos (for rpy_s / rpy_c)
void host_sincos(float x, float *s, float *c);

} // namespace doapp
