/*
 * doapp_oracle.h — CPU restatement of the reference's perturbator cost path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (doapp_b200/, include/)
 * includes, links or calls this; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs do, as the checker or the
 * timed CPU baseline.
 *
 * Each function cites the reference lines it restates (paths relative to the
 * reference tree, "gt.cu" = perturbator/src/generate_trajectories.cu,
 * "ps.cu" = perturbator/src/planning_state.cu, "cdg.cu" =
 * perturbator/src/cpu_distance_grid.cu, "og.cpp" =
 * occupancy_grid/src/occupancy_grid.cpp, "gt.test.cu" =
 * perturbator/src/generate_trajectories.test.cu).
 *
 * Parity pinning: the restatement is checked (tests/test_oracle_ref.py) against
 * the reference's own sources compiled for the host by oracle/build_ref.sh
 * into oracle/_ref/ (derivatives, perturbation, smoothness, scores, grid lookup,
 * Matrix ops, the 2x2 golden product of perturbator/src/test_matrix.cu).
 * The collision term has no runnable reference (it is stubbed at gt.cu:394 and
 * its mat-vec product is wrong): that stage is pinned only by
 * this restatement of gt.cu:397-463 -> "parity unpinned" for FK+collision.
 */
#ifndef DOAPP_ORACLE_H
#define DOAPP_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_COLLISION_AS_SHIPPED = 0, ORC_COLLISION_INTENDED = 1 };
enum { ORC_SELECT_CEIL_INT = 0, ORC_SELECT_FLOAT = 1 };
/* KEEP_BEST = the live path gt.cu:225-241; WEIGHTED = the commented-out update gt.cu:117-142,242-251 */
enum { ORC_UPDATE_KEEP_BEST = 0, ORC_UPDATE_WEIGHTED = 1 };
enum { ORC_GRID_NONE = 0, ORC_GRID_U8 = 1, ORC_GRID_F32 = 2 };
enum { ORC_FRAME_CDG = 0, ORC_FRAME_OG = 1 };
/* order of the sum over waypoints of the per-waypoint sum of acc^2:
 * SEQUENTIAL = the host loop of gt.test.cu:187-195;
 * TREE32 = lane-strided partials + xor butterfly (the fixed order the CUDA
 * kernels use in place of the reference's unordered float atomicAdd, gt.cu:73-77). */
/* SEG4 = left folds inside segments of 4 waypoints, then a left fold of the segment partials.
 * WAYPOINT (default) = per-waypoint sums over the joints (joint 0 first), then one left fold
 * over the waypoints (waypoint 0 first): the fixed order of the CUDA kernels, independent of
 * how many threads share a sample.  It differs from SEQUENTIAL only in that the d squares of
 * a waypoint are summed before they join the running total. */
enum { ORC_SUM_SEQUENTIAL = 0, ORC_SUM_TREE32 = 1, ORC_SUM_SEG4 = 2, ORC_SUM_WAYPOINT = 3 };

typedef struct orc_grid {
  int32_t kind;  /* ORC_GRID_* */
  int32_t frame; /* ORC_FRAME_* */
  uint32_t nx, ny, nz;
  double resolution;
  void *cells; /* uint8_t* or float*, dense ix + nx*iy + nx*ny*iz */
} orc_grid;

typedef struct orc_params {
  uint32_t k, m, n, d, n_noise;
  float delta_t, smooth_scale, collision_distance;
  int32_t collision_mode, select_mode, sum_order;
  float limit_weight;
  const float *limit_min, *limit_max; /* [d] or NULL */
  float sigma;
  uint64_t seed;
  uint64_t sample_offset, m_global;
  int32_t update_mode; /* ORC_UPDATE_* */
  float update_gain;   /* 10000.0f, gt.cu:140 */
  const struct orc_fk_model *fk_model; /* NULL: the chain of gt.cu:397-463; else SURVEY 8f row 3 */
} orc_params;

/* URDF-correct forward kinematics with per-link collision spheres:
 * accumulated transforms T_i = T_{i-1} * Trans(origin_i) * Rot(axis_i, theta_i), joint origins
 * and axes from rexarm_description/robots/rexarm.urdf:61-66,80-85,106-111,125-130,144-149;
 * spheres are fixed points of the link frames placed along the link boxes (:49-58,68-77,
 * 94-103,113-122,132-141,174-183,194-198).  A sphere collides when the occupancy cell at its
 * centre is set (U8 grid) or when the distance cell there is <= radius + collision_distance. */
#define ORC_FK_MAX_SPHERES 16
typedef struct orc_fk_sphere {
  int32_t link;    /* frame after joint `link`, 0..4 */
  float centre[3]; /* in that frame */
  float radius;
} orc_fk_sphere;
typedef struct orc_fk_model {
  float joint_origin[5][3];
  int32_t joint_axis[5]; /* 0 = x, 1 = y, 2 = z */
  uint32_t num_spheres;
  orc_fk_sphere spheres[ORC_FK_MAX_SPHERES];
} orc_fk_model;
void orc_fk_model_rexarm(orc_fk_model *m);
/* out[num_spheres][3]; a 4-joint waypoint has wrist_rotate = 0 */
void orc_fk_model_points(const orc_fk_model *m, const float *waypoint, uint32_t d, float (*out)[3]);

void orc_params_init(orc_params *p);

/* deterministic elementary functions (DESIGN.md "deterministic math spec") */
void orc_sincosf(float x, float *s, float *c);
float orc_logf(float x);

/* Philox4x32-10 (Salmon et al., SC'11; Random123 v1.09 known answers) */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
/* noise[k][m][n_noise][d] for one iteration of this shard */
void orc_noise_fill(const orc_params *p, uint32_t iteration, float *noise);
/* one element (global sample index t_global of climber c) */
float orc_noise_at(const orc_params *p, uint32_t iteration, uint32_t c, uint64_t t_global,
                   uint32_t wn, uint32_t j);

/* gt.cu:397-463 with a true mat-vec product: 5 link points of one waypoint */
void orc_rotation(int axis, float theta, float out[9]);
void orc_rotation_rpy(float roll, float pitch, float yaw, float out[9]);
void orc_fk_points(const float *waypoint, uint32_t d, float out[5][3]);
/* same, but with operator*(Matrix,Vector) as shipped (matrix.cuh:315-336: row sums) */
void orc_fk_points_as_shipped_matvec(const float *waypoint, uint32_t d, float out[5][3]);

/* cdg.cu:113-126 / og.cpp:29-40: flat cell index or -1 when outside */
int64_t orc_grid_index(const orc_grid *g, float x, float y, float z);
int orc_grid_collides(const orc_grid *g, float x, float y, float z, float collision_distance);
/* og.cpp:42-58 applied to every point of a cloud; returns #points inside */
size_t orc_grid_scatter(orc_grid *g, const float *xyz, size_t num_points, int8_t val,
                        int clear_first);
/* mapping/src/mapping_node.cpp:89-137 (deprojection of uint16 millimetre depth, 0 = invalid),
 * :195 (rigid transform into the arm frame; products summed left to right, then the
 * translation -- PCL/Eigen's own order is third-party and absent: unpinned), :262-263 (crop
 * box, inclusive), then og.cpp:47-58 per point.  depth [cams][h][w]; intrinsics [cams][4]
 * fx,fy,cx,cy; extrinsics [cams][12] r00..r22,tx,ty,tz.  Returns the points that passed the crop. */
uint64_t orc_depth_scatter(orc_grid *g, const uint16_t *depth, uint32_t cams, uint32_t w, uint32_t h,
                           const double *intrinsics, const float *extrinsics, double crop_size,
                           int8_t val, int clear_first);
/* Occupancy -> exact Euclidean distance field ("next" row 2 of SURVEY 8f; replaces the host
 * KD-tree fill CPUDistanceGrid::update, cdg.cu:87-202, whose per-voxel nearest-neighbour
 * distance this computes for points = centres of the occupied voxels).  Squared distances are
 * exact integers in voxel units (separable lower-envelope scan, Meijster et al. 2000);
 * out[v] = sqrtf((float)d2) * resolution, +inf when the grid is empty.  occ is a U8 grid. */
void orc_distance_field(const orc_grid *occ, float *out);
/* the same by exhaustive search (small grids only): the check of the scan above */
void orc_distance_field_brute(const orc_grid *occ, float *out);
/* gt.cu:391-466 */
uint32_t orc_num_collisions(const orc_params *p, const orc_grid *g, const float *waypoint);

/* gt.cu:46-59 + :276-291 + :61-115 for one sample.
 * base [n][d], start [d], noise [n_noise][d].  Optional outputs (may be NULL):
 * noisy/vel/acc [n][d], smooth, violated (1 if any joint outside limits). */
float orc_cost_sample(const orc_params *p, const orc_grid *g, const float *base,
                      const float *start, const float *noise, int32_t *collisions,
                      float *noisy, float *vel, float *acc, float *smooth, int32_t *violated);

/* every sample of every climber + host argmin (ps.cu:113-122).
 * base_traj [k][n][d]; noise [k][m][n_noise][d] or NULL -> Philox(iteration).
 * scores/collisions [k][m].  Returns flat argmin c*m_global + t_global (-1 none). */
int64_t orc_cost_eval(const orc_params *p, const orc_grid *g, const float *base_traj,
                      const float *start, const float *noise, uint32_t iteration,
                      float *scores, int32_t *collisions, float *min_score, int num_threads);

/* ps.cu:46-54 with the missing write restored */
void orc_plan_lerp(const float *start, const float *goal, uint32_t n, uint32_t d, float *out);

/* gt.cu:144-258 (keep-best loop, D5/D6 fixed) for `iters` iterations; noise
 * NULL -> Philox, else [iters][k][m][n_noise][d].  Outputs as doapp_plan. */
void orc_plan(const orc_params *p, const orc_grid *g, const float *start, const float *init_traj,
              uint32_t iters, const float *noise, float *best_traj, float *best_vel,
              float *best_score, int64_t *best_climber, float *climber_traj,
              float *climber_scores, int32_t *climber_rounded, uint8_t *found_better,
              int num_threads);

/* gt.cu:117-142 / gt.test.cu:200-208: u[c][j] = sum_t (best/ceil(score_t)) * noise[t][j] */
void orc_update_vector(const orc_params *p, const float *scores, const float *noise,
                       float *update /*[k][d]*/);

/* gt.cu:117-133 for ONE climber with an order-independent sum (the reference's float
 * atomicAdd order is arbitrary): every term weight*noise is rounded to a
 * multiple of 2^-28 and the terms are summed as 64-bit integers, so the result does not
 * depend on the order, the launch geometry or the number of GPUs.
 *   best_rounded  the block's atomicMin of int(ceilf(score))          (gt.cu:109-111)
 *   scores [m], noise [m][d] (n_noise == 1)
 *   sums[j] += round_to_even(term * 2^28);  sums[7] counts terms that are not finite or
 *   >= 4096 in magnitude ("poisoned": the update vector becomes NaN). */
void orc_update_sums(const orc_params *p, int32_t best_rounded, const float *scores,
                     const float *noise, uint32_t count, int64_t sums[8]);
/* u[j] from the integer sums; NaN if poisoned */
float orc_update_from_sum(int64_t sum, int64_t poisoned);

/* ps.cu:179-209 (D11 fixed): (n+2) points of positions/velocities [n+2][d], times [n+2] */
void orc_build_message(const float *cur, const float *goal, const float *traj, const float *vel,
                       uint32_t n, uint32_t d, float delta_t, float *positions,
                       float *velocities, float *time_from_start);

#ifdef __cplusplus
}
#endif
#endif
