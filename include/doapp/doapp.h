/*
 * doapp.h — C ABI of the B200-native perturbator hot path.
 *
 * This is the drop-in boundary.  Every entry point below
 * replaces one piece of the reference's `PlanningState` / `optimize_trajectories`
 * / `CPUDistanceGrid` / `OccupancyGrid` surface; the reference location is cited
 * next to each declaration (paths relative to the reference tree).  The C++
 * class `PlanningState` in doapp_b200/host/planning_state.hpp is implemented
 * purely on top of these calls, and INTEGRATION.md shows the binding a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - plain C, plain pointers and sizes; no C++/torch types cross the ABI.
 *   - pointers are HOST pointers unless the function name ends in `_dev`.
 *   - nothing aborts or throws: every call returns a doapp_status and leaves a
 *     message retrievable with doapp_last_error() (the reference aborts with
 *     ROS_ASSERT(false) on any CUDA error, perturbator/src/planning_state.cu:37-43).
 *   - a ctx is not thread-safe; distinct ctxs are independent.
 *   - there is NO CPU fallback: without a CUDA device doapp_create() fails.
 *
 * Symbols (reference notation, perturbator/src/main.cu:20):
 *   k = climbers (initial trajectories, one CUDA block each in the reference)
 *   m = samples (noise vectors) per climber, S = k*m samples per iteration
 *   n = waypoints per trajectory, d = joints per waypoint.
 */
#ifndef DOAPP_DOAPP_H
#define DOAPP_DOAPP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define DOAPP_API
#else
#define DOAPP_API __attribute__((visibility("default")))
#endif

#define DOAPP_ABI_VERSION 1

typedef struct doapp_ctx doapp_ctx;

typedef enum doapp_status {
  DOAPP_OK = 0,
  DOAPP_INVALID_ARG = 1,
  DOAPP_CUDA_ERROR = 2,
  DOAPP_NCCL_ERROR = 3,
  DOAPP_OOM = 4,
  DOAPP_NOT_READY = 5
} doapp_status;

/* Collision term.  AS_SHIPPED reproduces the `return 0; //Quick hack` at
 * perturbator/src/generate_trajectories.cu:394 (collision cost == 0);
 * INTENDED evaluates the FK chain + 5 grid lookups of :397-463. */
typedef enum doapp_collision_mode {
  DOAPP_COLLISION_AS_SHIPPED = 0,
  DOAPP_COLLISION_INTENDED = 1
} doapp_collision_mode;

/* How keep-best compares scores.  CEIL_INT is the kernel's
 * `int(ceilf(score))` + atomicMin (generate_trajectories.cu:109-111,225-228);
 * FLOAT is the host argmin's strict `<` (planning_state.cu:116-122). */
typedef enum doapp_select_mode {
  DOAPP_SELECT_CEIL_INT = 0,
  DOAPP_SELECT_FLOAT = 1
} doapp_select_mode;

/* Per-iteration update rule.  KEEP_BEST is the live code path
 * (generate_trajectories.cu:225-241, with the reversed assignment fixed);
 * WEIGHTED is the commented-out update of :117-142,242-251. */
typedef enum doapp_update_mode {
  DOAPP_UPDATE_KEEP_BEST = 0,
  DOAPP_UPDATE_WEIGHTED = 1
} doapp_update_mode;

/* Voxel grid element type. */
typedef enum doapp_grid_kind {
  DOAPP_GRID_NONE = 0,
  DOAPP_GRID_U8 = 1, /* occupancy, collision iff cell != 0 (occupancy_grid.cpp) */
  DOAPP_GRID_F32 = 2 /* distance, collision iff cell <= collision_distance
                        (generate_trajectories.cu:389,412) */
} doapp_grid_kind;

/* World frame -> voxel index convention.
 * CDG: perturbator/src/cpu_distance_grid.cu:69-82,113-126 — x,y centred,
 *      z in [0, H*r), fp32 add + IEEE fp32 divide + truncation.
 * OG : occupancy_grid/src/occupancy_grid.cpp:29-40 — cube centred on all three
 *      axes, fp64 `(c + size/2) / granularity`, truncation. */
typedef enum doapp_grid_frame {
  DOAPP_FRAME_CDG = 0,
  DOAPP_FRAME_OG = 1
} doapp_grid_frame;

/* Construction parameters: the literals the reference hard-codes in
 * perturbator/src/main.cu:18-21, planning_state.cuh:55-56 and
 * generate_trajectories.cu:23,389 gathered into one POD. */
typedef struct doapp_config {
  uint32_t struct_size;         /* = sizeof(doapp_config) */
  int32_t device;               /* CUDA device ordinal */
  uint32_t num_climbers;        /* k */
  uint32_t samples_per_climber; /* m handled by THIS ctx (local shard) */
  uint32_t num_waypoints;       /* n, 1..1024 */
  uint32_t waypoint_dim;        /* d, 4 or 5 */
  uint32_t noise_waypoints;     /* 1 (reference: one d-vector per sample,
                                   generate_trajectories.cu:37-59) or n */
  float delta_t;                /* 0.1f  (planning_state.cuh:55) */
  float smooth_scale;           /* 0.25f (two `*= 0.5f`, generate_trajectories.cu:81,107) */
  float collision_distance;     /* 0.05f (generate_trajectories.cu:389) */
  int32_t collision_mode;       /* doapp_collision_mode */
  int32_t select_mode;          /* doapp_select_mode */
  int32_t update_mode;          /* doapp_update_mode */
  float update_gain;            /* 10000.0f (generate_trajectories.cu:140) */
  float limit_weight;           /* weight of the joint-limit hinge cost; 0 = off
                                   (the reference only has a TODO, :113-114) */
  float sigma;                  /* std-dev of generated noise; reference = 1.0f */
  uint64_t seed;                /* 1234 (generate_trajectories.cu:23) */
  /* sharding over ranks: this ctx owns samples
   * [sample_offset, sample_offset + samples_per_climber) of each climber's
   * global_samples_per_climber; sample identity (noise counter, argmin index)
   * is the GLOBAL index so results do not depend on the number of ranks. */
  uint64_t sample_offset;
  uint64_t global_samples_per_climber; /* 0 -> samples_per_climber */
} doapp_config;

/* Fill *cfg with the reference defaults listed above (k=1, m=1024, n=64, d=5). */
DOAPP_API void doapp_config_init(doapp_config *cfg);

DOAPP_API uint32_t doapp_version(void); /* DOAPP_ABI_VERSION */
DOAPP_API const char *doapp_status_str(doapp_status s);

/* PlanningState::PlanningState (planning_state.cu:9-44): allocates every
 * device buffer once (plain cudaMalloc, not managed memory). */
DOAPP_API doapp_status doapp_create(const doapp_config *cfg, doapp_ctx **out);
DOAPP_API void doapp_destroy(doapp_ctx *ctx);
/* Message of the last failing call on this ctx ("" if none); ctx may be NULL
 * to read the message of a failed doapp_create on this thread. */
DOAPP_API const char *doapp_last_error(const doapp_ctx *ctx);

/* ---- world model (L1) ------------------------------------------------- */

/* CPUDistanceGrid ctor (cpu_distance_grid.cu:28-83, Dimensions at
 * cpu_distance_grid.cuh:13-18) / OccupancyGrid ctor (occupancy_grid.cpp:8-17):
 * allocate a dense nx*ny*nz grid of `kind`, zero-filled for U8 and +inf for
 * F32.  For DOAPP_FRAME_OG nx==ny==nz and resolution is the granularity. */
DOAPP_API doapp_status doapp_grid_alloc(doapp_ctx *ctx, int32_t kind, int32_t frame,
                                        uint32_t nx, uint32_t ny, uint32_t nz,
                                        double resolution);
/* Copy a dense host grid (index ix + nx*iy + nx*ny*iz) to the device. */
DOAPP_API doapp_status doapp_grid_upload(doapp_ctx *ctx, const void *host_cells,
                                         size_t num_cells);
DOAPP_API doapp_status doapp_grid_download(doapp_ctx *ctx, void *host_cells,
                                           size_t num_cells);
/* OccupancyGrid::deserialize (occupancy_grid.cpp:107-113) with the three
 * fields of occupancy_grid/msg/OccupancyGrid.msg:1-3.  (Re)allocates a U8 /
 * FRAME_OG grid with N = size_t(size/granularity) cells per edge. */
DOAPP_API doapp_status doapp_grid_from_msg(doapp_ctx *ctx, double size, double granularity,
                                           const int8_t *grid, size_t len);
/* Stage 5: OccupancyGrid::set(double,double,double,val) applied to every point
 * of an xyz cloud (occupancy_grid.cpp:47-58; the reference has no such loop,
 * SURVEY D13).  Points outside the grid are skipped.  U8 grids only. */
DOAPP_API doapp_status doapp_grid_scatter(doapp_ctx *ctx, const float *xyz, size_t num_points,
                                          int8_t val, int clear_first);
DOAPP_API doapp_status doapp_grid_scatter_dev(doapp_ctx *ctx, const float *xyz_dev,
                                              size_t num_points, int8_t val, int clear_first);
/* Depth frames -> occupancy in one pass ("next" row of SURVEY 8f): the mapping node's
 * deprojection (mapping/src/mapping_node.cpp:89-137: uint16 millimetres, 0 = invalid,
 * x = (u - cx) * depth * (0.001 / fx)), its transform into the arm frame (:195, extrinsics
 * :63-82 given here as a row-major rotation + translation, camera -> arm_bundle), its crop box
 * (:262-263: |x|,|y| <= crop_size/2, 0 <= z <= crop_size) and OccupancyGrid::set per surviving
 * point (occupancy_grid.cpp:47-58), fused so that no point cloud is materialised.
 *   depth      [num_cams][height][width] uint16
 *   intrinsics [num_cams][4] double  fx, fy, cx, cy   (image_geometry::PinholeCameraModel)
 *   extrinsics [num_cams][12] float  r00..r22, tx, ty, tz
 *   points_kept out (may be NULL): points that passed the crop box
 * U8 grids only; 1..4 cameras. */
DOAPP_API doapp_status doapp_grid_from_depth(doapp_ctx *ctx, const uint16_t *depth,
                                             uint32_t num_cams, uint32_t width, uint32_t height,
                                             const double *intrinsics, const float *extrinsics,
                                             double crop_size, int8_t val, int clear_first,
                                             uint64_t *points_kept);
/* Occupancy -> distance field ("next" row of SURVEY 8f): replaces the host KD-tree fill
 * CPUDistanceGrid::update (cpu_distance_grid.cu:87-202, one nanoflann query per voxel on the
 * host) by an exact Euclidean distance transform on the device.  The ctx's U8 grid becomes an
 * F32 grid of the same frame and dimensions whose cell v holds the distance from v's centre to
 * the centre of the nearest occupied voxel (+inf if none), so that the collision predicate is
 * the reference's `distance <= collision_distance` (generate_trajectories.cu:389,412).
 * Every dimension must be <= 768. */
DOAPP_API doapp_status doapp_grid_distance_field(doapp_ctx *ctx);
/* Occupancy -> exact Euclidean distance field WITHOUT any allocation.  No counterpart in the
 * reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same three launches and the same bits as doapp_grid_distance_field, but the transform's
 * scratch (two int32 planes) is a static device array inside the library, so the entry point
 * calls no cudaMalloc / cudaFree and needs no doapp_ctx.  Every GPU that loads the library
 * holds DOAPP_NOALLOC_MAX_CELLS * 8 bytes = 16 MiB for it until the process exits.
 *   occ_dev   DEVICE pointer, nx*ny*nz uint8 occupancy (non-zero = occupied), index
 *             ix + nx*iy + nx*ny*iz, on the CURRENT device
 *   dist_dev  DEVICE pointer, nx*ny*nz fp32: distance to the nearest occupied voxel centre in
 *             metres (voxels * resolution), +inf for an empty grid
 *   stream    a cudaStream_t (NULL = the default stream); the call only enqueues, the caller
 *             synchronises
 * Grids above DOAPP_NOALLOC_MAX_CELLS cells (128^3) or 768 cells per edge are rejected with
 * DOAPP_INVALID_ARG and a message in doapp_last_error(NULL).  ALL calls on a device share the
 * one scratch array: it keeps the last call's contents (harmless: every cell is written before
 * it is read), and two calls that overlap on the GPU (different streams) RACE -- serialise
 * them.  The first call on a device resolves the array's address (cudaGetSymbolAddress): make
 * it once outside any stream capture. */
#define DOAPP_NOALLOC_MAX_CELLS (128ull * 128ull * 128ull)
DOAPP_API doapp_status doapp_distance_field_noalloc(const uint8_t *occ_dev, uint32_t nx,
                                                    uint32_t ny, uint32_t nz, double resolution,
                                                    float *dist_dev, void *stream);
/* CPUDistanceGrid::operator() / OccupancyGrid::get(double...) on a batch of
 * points: writes the flat cell index (-1 if outside) and the collision
 * predicate for each point.  Debug / parity helper. */
DOAPP_API doapp_status doapp_grid_lookup(doapp_ctx *ctx, const float *xyz, size_t num_points,
                                         int64_t *cell_index, uint8_t *collides);

/* ---- cost inputs ------------------------------------------------------ */

/* min_joint_angles / max_joint_angles (perturbator/src/common.cu:7-15). */
DOAPP_API doapp_status doapp_set_limits(doapp_ctx *ctx, const float *min_d, const float *max_d);

/* ---- URDF-correct forward kinematics with per-link collision spheres ---------------
 * The reference's chain (generate_trajectories.cu:397-463) rotates a point base->tip about
 * y,z,y,y,y; the arm's description says z,y,y,y,z with accumulated link frames
 * (rexarm_description/robots/rexarm.urdf:61-66,80-85,106-111,125-130,144-149).
 * With a model set, the collision term of every cost evaluation walks the URDF chain
 *   p_i = p_{i-1} + R_{i-1} * origin_i,   R_i = R_{i-1} * Rot(axis_i, theta_i)
 * and counts the spheres (fixed points of the link frames, sized from the link boxes of
 * rexarm.urdf:49-58,68-77,94-103,113-122,132-141,174-183) that collide: occupancy cell at the
 * centre != 0, or distance cell at the centre <= radius + collision_distance. */
#define DOAPP_FK_MAX_SPHERES 16
typedef struct doapp_fk_sphere {
  int32_t link;    /* frame after joint `link`, 0..4 */
  float centre[3]; /* in that frame, metres */
  float radius;
} doapp_fk_sphere;
typedef struct doapp_fk_model {
  uint32_t struct_size;      /* = sizeof(doapp_fk_model) */
  float joint_origin[5][3];  /* rexarm.urdf:64,83,109,128,147 */
  int32_t joint_axis[5];     /* 0 = x, 1 = y, 2 = z; rexarm.urdf:65,84,110,129,148 */
  uint32_t num_spheres;      /* 1..DOAPP_FK_MAX_SPHERES */
  doapp_fk_sphere spheres[DOAPP_FK_MAX_SPHERES];
} doapp_fk_model;
/* The Rexarm of rexarm.urdf: 5 joints, two spheres along every link box, two along the
 * fingers, one at gripper_center (13 spheres). */
DOAPP_API void doapp_fk_model_rexarm(doapp_fk_model *model);
/* model == NULL returns to the reference's chain (the default). */
DOAPP_API doapp_status doapp_set_fk_model(doapp_ctx *ctx, const doapp_fk_model *model);

/* ---- stage 1: noise --------------------------------------------------- */

/* Philox4x32-10 + Box-Muller replacement for init_cudarand / curand_normal
 * (generate_trajectories.cu:17-25,37-44,297-299).  Writes the noise the fused
 * kernels would generate for `iteration` to host memory, layout
 * [k][m][noise_waypoints][d] for this ctx's shard. */
DOAPP_API doapp_status doapp_noise_generate(doapp_ctx *ctx, uint32_t iteration, float *noise_out);

/* ---- stages 2-4: one cost-evaluation pass ------------------------------ */

/* score_noisy_trajectories for every sample (generate_trajectories.cu:46-115)
 * followed by the host argmin (planning_state.cu:113-122).
 *   base_traj [k][n][d]   current trajectory of every climber
 *   start     [d]         initial_waypoint (generate_trajectories.cu:10)
 *   noise     [k][m][noise_waypoints][d], or NULL -> Philox(seed, iteration)
 *   scores    [k][m]      out, may be NULL
 *   collisions[k][m]      out, may be NULL
 *   argmin                out, flat index c*m_global + t_global of the first
 *                         lowest score (strict <), -1 if no finite score
 *   min_score             out */
DOAPP_API doapp_status doapp_cost_eval(doapp_ctx *ctx, const float *base_traj, const float *start,
                                       const float *noise, uint32_t iteration, float *scores,
                                       int32_t *collisions, int64_t *argmin, float *min_score);

/* Device-resident variant used by the bench: inputs were staged with
 * doapp_stage_inputs(); nothing crosses PCIe.  Runs `repeats` passes
 * back to back on the ctx stream and returns after synchronising. */
DOAPP_API doapp_status doapp_stage_inputs(doapp_ctx *ctx, const float *base_traj,
                                          const float *start, const float *noise);
DOAPP_API doapp_status doapp_cost_eval_dev(doapp_ctx *ctx, uint32_t iteration, int use_staged_noise,
                                           uint32_t repeats, float *elapsed_ms);
DOAPP_API doapp_status doapp_read_results(doapp_ctx *ctx, float *scores, int32_t *collisions,
                                          int64_t *argmin, float *min_score);

/* ---- the planner loop -------------------------------------------------- */

/* PlanningState::plan_lerp with the missing write restored (planning_state.cu:46-54,
 * SURVEY D4): out[w][j] = start[j] + (w+1)*(goal[j]-start[j])/n. Host only. */
DOAPP_API doapp_status doapp_plan_lerp(const float *start, const float *goal, uint32_t n,
                                       uint32_t d, float *out);

/* PlanningState::plan() + publish_trajectory()'s argmin (planning_state.cu:56-90,
 * 108-127) with a bounded iteration count instead of the host-polled stop flag
 * (generate_trajectories.cu:253).  Every climber starts from `init_traj`
 * ([n][d], normally doapp_plan_lerp) and runs `iters` iterations of
 * noise -> cost -> argmin -> update.
 *   noise      NULL -> Philox; else [iters][k][m_global][noise_waypoints][d]
 *   best_traj  [n][d] out;  best_vel [n][d] out (backward differences)
 *   best_score out;  best_climber out (-1 -> no climber improved on init_traj) */
DOAPP_API doapp_status doapp_plan(doapp_ctx *ctx, const float *start, const float *goal,
                                  const float *init_traj, uint32_t iters, const float *noise,
                                  float *best_traj, float *best_vel, float *best_score,
                                  int64_t *best_climber);
/* plan() / publish_trajectory() split (planning_state.cu:79-84 launches, :111
 * blocks): _async enqueues the whole loop and returns, _wait blocks and reads
 * the result back. */
DOAPP_API doapp_status doapp_plan_async(doapp_ctx *ctx, const float *start, const float *goal,
                                        const float *init_traj, uint32_t iters,
                                        const float *noise);
DOAPP_API doapp_status doapp_plan_wait(doapp_ctx *ctx, float *best_traj, float *best_vel,
                                       float *best_score, int64_t *best_climber);
/* Per-climber state after a plan: trajectories [k][n][d], float score and the
 * rounded score of each climber's kept trajectory, found_better flags
 * (generate_trajectories.cu:254-257). Any pointer may be NULL. */
DOAPP_API doapp_status doapp_plan_state(doapp_ctx *ctx, float *trajectories, float *scores,
                                        int32_t *rounded, uint8_t *found_better);
/* Dynamic-obstacle replan (BASELINE config 3): every iteration first rebuilds
 * the grid from clouds[it] ([iters][num_points][3]), then runs one planner
 * iteration.  The clouds are copied in one by one on a second stream; iteration
 * it waits for cloud it only, so the copy of the next cloud overlaps the replan.
 * clouds == NULL reuses the clouds the previous call left in device memory
 * (same iters and num_points).  elapsed_ms: CUDA events around the loop. */
DOAPP_API doapp_status doapp_replan_dynamic(doapp_ctx *ctx, const float *start, const float *goal,
                                            const float *init_traj, uint32_t iters,
                                            const float *clouds, size_t num_points,
                                            float *best_traj, float *best_score,
                                            float *elapsed_ms);
/* Bench hook: re-run the last planned loop `repeats` times from the staged
 * inputs, timing with CUDA events on the ctx stream. */
DOAPP_API doapp_status doapp_plan_bench(doapp_ctx *ctx, uint32_t iters, uint32_t repeats,
                                        float *elapsed_ms);

/* Profiling hook: one un-graphed pass of the staged planner loop with a CUDA
 * event pair around every cost-kernel launch on the ctx stream; returns the
 * summed cost-kernel time, the number of launches (iters, or 1 when the whole loop
 * runs as one persistent cooperative launch) and the whole loop's time. */
DOAPP_API doapp_status doapp_plan_profile(doapp_ctx *ctx, uint32_t iters, float *cost_ms_total,
                                          uint32_t *cost_launches, float *loop_ms);

/* Roofline probe: rate of independent random single-cell loads (one 32-byte
 * sector each) from a scratch array of `array_bytes`, the memory-side ceiling of the
 * collision lookups of generate_trajectories.cu:412 / cpu_distance_grid.cu:113-126.  An array
 * well inside L2 gives the L2 gather rate, one larger than L2 the HBM rate. */
DOAPP_API doapp_status doapp_probe_gather(doapp_ctx *ctx, size_t array_bytes, uint32_t cell_bytes,
                                          double *gathers_per_sec);

/* ---- multi-GPU ------------------------------------------- */

#define DOAPP_COMM_ID_BYTES 128
/* ncclGetUniqueId; rank 0 calls it and ships the bytes to its peers. */
DOAPP_API doapp_status doapp_comm_unique_id(uint8_t id[DOAPP_COMM_ID_BYTES]);
/* ncclCommInitRank on the ctx's device.  After this, doapp_plan* merges the
 * per-iteration argmin keys of all ranks with one 8*k-byte min all-reduce, and
 * doapp_grid_broadcast replicates rank `root`'s grid. */
DOAPP_API doapp_status doapp_comm_init(doapp_ctx *ctx, const uint8_t id[DOAPP_COMM_ID_BYTES],
                                       int32_t rank, int32_t nranks);
DOAPP_API doapp_status doapp_grid_broadcast(doapp_ctx *ctx, int32_t root);

/* Peer-memory argmin merge (NVLink P2P).  With the ranks' mailboxes mapped into each other,
 * the persistent planner kernel trades the per-iteration argmin keys itself -- one 8-byte
 * store per peer and iteration, inside the kernel -- instead of one NCCL all-reduce launch
 * per iteration.  Each rank exports a CUDA IPC handle of its mailbox, the caller gathers
 * the handles of all ranks (any transport; same role as the NCCL unique id above) and every
 * rank imports the full table, then all ranks synchronise once on the host before the first
 * doapp_plan.  At most 8 ranks (one NVSwitch domain).  Optional: without it multi-rank plans
 * use the NCCL path; doapp_peer_import(ctx, NULL, 0, 0) switches back to it.
 * The ranks' kernels wait for each other: drive the ranks from separate threads / processes, or
 * call doapp_plan_async on every rank's ctx before the first doapp_plan_wait.  A rank that never
 * arrives makes the others give up after 20 s (doapp_plan_wait returns DOAPP_CUDA_ERROR). */
#define DOAPP_PEER_HANDLE_BYTES 64
DOAPP_API doapp_status doapp_peer_export(doapp_ctx *ctx, uint8_t handle[DOAPP_PEER_HANDLE_BYTES]);
DOAPP_API doapp_status doapp_peer_import(doapp_ctx *ctx, const uint8_t *handles /*[nranks][64]*/,
                                         int32_t rank, int32_t nranks);

/* ---- introspection ----------------------------------------------------- */

/* How the last planner loop ran: bit 0 = one persistent cooperative launch (else one cost
 * launch per iteration), bit 1 = argmin keys merged in-kernel through the peer mailboxes
 * (else, with several ranks, one NCCL all-reduce per iteration). */
DOAPP_API uint32_t doapp_plan_mode(const doapp_ctx *ctx);
/* Counters of the last plan when it ran as one persistent launch (else zeros): accepted
 * keep-best updates (generate_trajectories.cu:225-241) and cost passes the first CTA made --
 * passes beyond `iters` were speculative ones discarded after an accepted update. */
DOAPP_API doapp_status doapp_plan_stats(doapp_ctx *ctx, uint64_t *accepted_updates,
                                        uint64_t *cost_passes);
/* Number of kernels this ctx has launched since creation (bench `gpu_launches`). */
DOAPP_API uint64_t doapp_launch_count(const doapp_ctx *ctx);
/* The deterministic sin/cos used by the FK kernels, evaluated ON THE DEVICE for
 * a batch of angles (parity test against the oracle's restatement). */
DOAPP_API doapp_status doapp_debug_sincos(doapp_ctx *ctx, const float *angles, size_t count,
                                          float *sin_out, float *cos_out);
/* FK link points of generate_trajectories.cu:397-463 for a batch of waypoints
 * ([count][d] in, [count][5][3] out), evaluated on the device. */
DOAPP_API doapp_status doapp_debug_fk(doapp_ctx *ctx, const float *waypoints, size_t count,
                                      float *points_out);
/* Sphere centres of the FK model set with doapp_set_fk_model ([count][d] in,
 * [count][num_spheres][3] out, the caller's sphere order), evaluated on the device. */
DOAPP_API doapp_status doapp_debug_fk_model(doapp_ctx *ctx, const float *waypoints, size_t count,
                                            float *centres_out);

#ifdef __cplusplus
}
#endif
#endif /* DOAPP_DOAPP_H */
