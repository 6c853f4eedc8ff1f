// kernels.cu — hand-written sm_100a kernels of the perturbator hot path.
//
// Stage map:
//   cost_kernel            a1 (Philox noise, fused) + a2 perturb + a3 derivatives + a4 FK +
//                          a5 grid gathers + a6 score + a7 per-climber argmin keys + the pending
//                          update of the previous launch applied while staging
//                          (generate_trajectories.cu:225-241 keep-best, D5/D6 fixed; or the
//                          weighted update of :117-142).  Thread = sample x run of waypoints,
//                          1..32 threads per sample (clusters + DSMEM when > 4).
//   cost_kernel_pk         the same pass, one thread per sample, two waypoints per step in
//                          packed fma.rn.f32x2 (FFMA2): half the issue slots for the fp32 work
//   plan_persistent_kernel the whole keep-best loop of :197-253 in ONE cooperative launch
//                          (small problems), argmin keys merged across GPUs through NVLink
//                          peer mailboxes inside the kernel
//   update_vector_kernel   a8 compute_update_vector (:117-133), order-independent integer sums
//   final_kernel           host argmin of planning_state.cu:113-127 + velocities
//   scatter_kernel         a10 point cloud -> voxel occupancy (occupancy_grid.cpp:42-58)
//   depth_scatter_kernel   depth frames -> voxel occupancy (mapping_node.cpp:89-137,195,262-263)
//
// The noisy trajectory, velocities and accelerations the reference materialises in three
// global arrays per iteration (generate_trajectories.cu:202-210) live only in registers.  Sums
// over waypoints use one fixed order (per-waypoint sums, left fold over the waypoints) instead
// of the reference's unordered float atomicAdd (:73-77), so scores and the argmin are
// reproducible bit for bit and independent of the launch geometry.
#include "kernels.cuh"

#include <cooperative_groups.h>

#include <cstdlib>
#include <cstring>

#include "device_math.cuh"

namespace doapp {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;

// ---------------------------------------------------------------- grid lookup
// Flat cell index of a world point or -1 (outside).  CDG: cpu_distance_grid.cu:113-126
// (fp32 add, IEEE divide, truncate); OG: occupancy_grid.cpp:29-40 (fp64).
// Half-open bounds and an explicit index check replace the reference's
// inclusive upper bound.
__device__ __forceinline__ long long grid_index(const GridDev &g, float x, float y, float z) {
  if (g.frame == 0) {
    if (!(x >= -g.x_off && x < g.x_off) || !(y >= -g.y_off && y < g.y_off) ||
        !(z >= 0.0f && z < g.z_max))
      return -1;
    float fx = dm::div_(dm::add_(x, g.x_off), g.res);
    float fy = dm::div_(dm::add_(y, g.y_off), g.res);
    float fz = dm::div_(z, g.res);
    uint32_t ix = __float2uint_rz(fx), iy = __float2uint_rz(fy), iz = __float2uint_rz(fz);
    if (ix >= g.nx || iy >= g.ny || iz >= g.nz) return -1;
    return (long long)((size_t)g.nx * g.ny * iz + (size_t)g.nx * iy + ix);
  } else {
    double xs = __dadd_rn((double)x, g.half_d), ys = __dadd_rn((double)y, g.half_d),
           zs = __dadd_rn((double)z, g.half_d);
    if (!(xs >= 0.0 && ys >= 0.0 && zs >= 0.0)) return -1;
    double fx = __ddiv_rn(xs, g.res_d), fy = __ddiv_rn(ys, g.res_d), fz = __ddiv_rn(zs, g.res_d);
    if (!(fx < (double)g.nx && fy < (double)g.ny && fz < (double)g.nz)) return -1;
    unsigned long long ix = __double2ull_rz(fx), iy = __double2ull_rz(fy),
                       iz = __double2ull_rz(fz);
    if (ix >= g.nx || iy >= g.ny || iz >= g.nz) return -1;
    return (long long)(ix + (unsigned long long)g.nx * iy + (unsigned long long)g.nx * g.ny * iz);
  }
}

__device__ __forceinline__ int grid_collides(const GridDev &g, float x, float y, float z,
                                             float coll_dist) {
  long long idx = grid_index(g, x, y, z);
  if (idx < 0) return 0;
  if (g.kind == 1) return __ldg((const uint8_t *)g.cells + idx) != 0;
  return __ldg((const float *)g.cells + idx) <= coll_dist; // generate_trajectories.cu:412
}

// ------------------------------------------------------------------------ FK
// generate_trajectories.cu:397-463 with a true mat-vec product.  Each link does
// p <- R_axis(theta) * (p + t); the rotation matrices of :320-352 contain only
// c, s, -s, 0 and 1, so the FMA chains of matrix.cuh:274-313 collapse to the
// two-term forms below (zero terms add +-0, which changes no value).
struct P3 {
  float x, y, z;
};
__device__ __forceinline__ P3 rot_y(float s, float c, P3 q) {
  P3 u;
  u.x = dm::fma_(s, q.z, dm::mul_(c, q.x));
  u.y = q.y;
  u.z = dm::fma_(c, q.z, dm::mul_(-s, q.x));
  return u;
}
__device__ __forceinline__ P3 rot_z(float s, float c, P3 q) {
  P3 u;
  u.x = dm::fma_(-s, q.y, dm::mul_(c, q.x));
  u.y = dm::fma_(c, q.y, dm::mul_(s, q.x));
  u.z = q.z;
  return u;
}

// FAST: straight-line small-argument sin/cos (dm::sincos_small), valid for |angle| < 2^17:
// the caller has checked the angles (dm::fine_angle) or redoes the waypoint through the
// general path.
template <int D, bool FAST>
__device__ __forceinline__ void fk_points(const float (&th)[D], float rpy_s, float rpy_c,
                                          P3 (&out)[5]) {
  float s, c;
  P3 p, q;
  auto sc = [&](float a) {
    if (FAST) {
      dm::sincos_small(a, &s, &c);
    } else {
      dm::sincos(a, &s, &c);
    }
  };
  // base: t = (0, -0.0015, 0.0715), roty(theta0)            gt.cu:405-410
  sc(th[0]);
  q = P3{0.0f, -0.0015f, 0.0715f};
  p = rot_y(s, c, q);
  out[0] = p;
  // shoulder: t = (0, 0, 0.045), rotz(theta1)                gt.cu:416-421
  sc(th[1]);
  q = P3{p.x, p.y, dm::add_(p.z, 0.045f)};
  p = rot_z(s, c, q);
  out[1] = p;
  // elbow: t = (0, 0, 0.1), roty(theta2)                     gt.cu:427-432
  sc(th[2]);
  q = P3{p.x, p.y, dm::add_(p.z, 0.1f)};
  p = rot_y(s, c, q);
  out[2] = p;
  // wrist: t = (0, 0, 0.1), roty(theta3)                     gt.cu:438-443
  sc(th[3]);
  q = P3{p.x, p.y, dm::add_(p.z, 0.1f)};
  p = rot_y(s, c, q);
  out[3] = p;
  // wrist rotate: t = (0, -0.0125, 0.08), roty(theta4)       gt.cu:449-454
  q = P3{p.x, dm::add_(p.y, -0.0125f), dm::add_(p.z, 0.08f)};
  if (D >= 5) {
    sc(th[D >= 5 ? 4 : 0]);
    p = rot_y(s, c, q);
  } else {
    p = q; // theta4 := 0 -> identity
  }
  // fixed Rrpy(0, -pi/2, 0) and t = (0, -0.005, 0.07)        gt.cu:456-459
  p = rot_y(rpy_s, rpy_c, p);
  p = P3{p.x, dm::add_(p.y, -0.005f), dm::add_(p.z, 0.07f)};
  out[4] = p;
}

__device__ __forceinline__ uint32_t key_hi(int select_mode, float score) {
  if (select_mode == 0) // int(ceilf(score)), generate_trajectories.cu:109, saturating
    return (score < 2147483648.0f) ? (uint32_t)(int)ceilf(score) : 0x7FFFFFFFu;
  return (score != score) ? 0xFFFFFFFFu : __float_as_uint(score);
}
__device__ __forceinline__ uint32_t key_hi_float(float score) {
  return (score != score) ? 0xFFFFFFFFu : __float_as_uint(score);
}

// ---------------------------------------------------------------- cost kernel
// One thread = one sample x one run of waypoint segments.  The 32 lanes of a
// warp are 32 consecutive samples of one climber walking the SAME waypoints, so
// the base-trajectory reads are warp-uniform shared-memory broadcasts, the
// neighbouring link points of the 32 perturbed copies gather from neighbouring
// voxels, and xprev / vprev stay in registers (no shuffles, no scratch arrays:
// the reference materialises noisy trajectories, velocities and accelerations
// in global memory, generate_trajectories.cu:202-210).
//
// Sum order (fixed, independent of the launch geometry; oracle ORC_SUM_WAYPOINT):
// the d squares of a waypoint are summed joint 0 first, then the per-waypoint
// sums are folded left to right, waypoint 0 first.  With T > 1 threads per
// sample each thread owns a run of consecutive waypoints and leaves its
// per-waypoint sums in the shared memory of the cluster's rank-0 CTA
// (distributed shared memory); one thread per sample folds them in order.
constexpr int kMaxCostThreads = 512;

// Grid modes the kernel is specialised for.
enum : int { GM_NONE = 0, GM_U8_CDG = 1, GM_F32_CDG = 2, GM_GENERIC = 3 };

// thr: a distance cell collides when it is <= thr (collision_distance for the reference's point
// links, radius + collision_distance for the spheres of an FK model)
template <int GM, bool FIDX>
__device__ __forceinline__ int collides(const CostParams &P, P3 p, float thr) {
  if (GM == GM_NONE) return 0;
  if (GM == GM_GENERIC) return grid_collides(P.grid, p.x, p.y, p.z, thr);
  const GridDev &g = P.grid;
  const float ax = dm::add_(p.x, g.x_off), ay = dm::add_(p.y, g.y_off);
  if (FIDX) {
    // cpu_distance_grid.cu:115-122 with half-open bounds.  Inside iff the bit
    // patterns of x + x_off, y + y_off, z + 0 are below the per-axis thresholds as unsigned
    // integers (GridDev::tx_bits): a negative sum or NaN has the top bits set, +0 absorbs
    // z == -0, and the threshold is the smallest dividend whose voxel index reaches n.
    const float az = dm::add_(p.z, 0.0f);
    const bool in = (__float_as_uint(ax) < g.tx_bits) & (__float_as_uint(ay) < g.ty_bits) &
                    (__float_as_uint(az) < g.tz_bits);
    // :120-122, IEEE quotients through the verified reciprocal path
    const uint32_t ix = __float2uint_rz(dm::div_fast_unguarded(ax, P.res_div));
    const uint32_t iy = __float2uint_rz(dm::div_fast_unguarded(ay, P.res_div));
    const uint32_t iz = __float2uint_rz(dm::div_fast_unguarded(az, P.res_div));
    const uint32_t lin = (iz * g.ny + iy) * g.nx + ix; // < 2^32 cells (grid_alloc)
    int hit = 0;
    if (GM == GM_U8_CDG) {
      if (in) hit = __ldg((const uint8_t *)g.cells + lin) != 0;
    } else {
      if (in) hit = __ldg((const float *)g.cells + lin) <= thr; // generate_trajectories.cu:412
    }
    return hit;
  }
  bool in = (p.x >= -g.x_off) & (p.x < g.x_off) & (p.y >= -g.y_off) & (p.y < g.y_off) &
            (p.z >= 0.0f) & (p.z < g.z_max);
  const float fx = dm::div_(ax, g.res), fy = dm::div_(ay, g.res), fz = dm::div_(p.z, g.res);
  const uint32_t ix = __float2uint_rz(fx), iy = __float2uint_rz(fy), iz = __float2uint_rz(fz);
  in = in & (ix < g.nx) & (iy < g.ny) & (iz < g.nz);
  const uint32_t lin = in ? (iz * g.ny + iy) * g.nx + ix : 0u;
  if (GM == GM_U8_CDG) return in & (__ldg((const uint8_t *)g.cells + lin) != 0);
  return in & (__ldg((const float *)g.cells + lin) <= thr);
}

// ---- URDF-correct chain with per-link spheres ---------------------
// Link frames accumulate as rexarm.urdf:61-66,80-85,106-111,125-130,144-149 define them:
//   p_i = p_{i-1} + R_{i-1} * origin_i,   R_i = R_{i-1} * Rot(axis_i, theta_i),
// a sphere of link i sits at p_i + R_i * centre.  R is kept as three columns; R * v is
// fma(col2, v.z, fma(col1, v.y, col0 * v.x)) per component; a rotation about axis a with cyclic
// successors (b, c) replaces col_b <- fma(s, col_c, c * col_b), col_c <- fma(c, col_c, -s * col_b)
// (DESIGN.md "FK model" fixes this operation order; the parity tests hold both sides to it).
__device__ __forceinline__ P3 frame_apply(const P3 &c0, const P3 &c1, const P3 &c2, const P3 &pos,
                                          float vx, float vy, float vz) {
  P3 o;
  o.x = dm::add_(pos.x, dm::fma_(c2.x, vz, dm::fma_(c1.x, vy, dm::mul_(c0.x, vx))));
  o.y = dm::add_(pos.y, dm::fma_(c2.y, vz, dm::fma_(c1.y, vy, dm::mul_(c0.y, vx))));
  o.z = dm::add_(pos.z, dm::fma_(c2.z, vz, dm::fma_(c1.z, vy, dm::mul_(c0.z, vx))));
  return o;
}
__device__ __forceinline__ void frame_rotate(float s, float c, P3 &b, P3 &cc) {
  const P3 nb{dm::fma_(s, cc.x, dm::mul_(c, b.x)), dm::fma_(s, cc.y, dm::mul_(c, b.y)),
              dm::fma_(s, cc.z, dm::mul_(c, b.z))};
  const P3 nc{dm::fma_(c, cc.x, dm::mul_(-s, b.x)), dm::fma_(c, cc.y, dm::mul_(-s, b.y)),
              dm::fma_(c, cc.z, dm::mul_(-s, b.z))};
  b = nb;
  cc = nc;
}
// visit(q, centre) for every sphere, in link order.  The model reads are warp-uniform (constant
// bank operands in the cost kernels, where the model is part of the parameter block).
template <int D, bool FAST, typename Visit>
__device__ __forceinline__ void fk_model_walk(const FkModelDev &M, const float (&th)[D],
                                              Visit &&visit) {
  P3 c0{1.0f, 0.0f, 0.0f}, c1{0.0f, 1.0f, 0.0f}, c2{0.0f, 0.0f, 1.0f}, pos{0.0f, 0.0f, 0.0f};
  uint32_t q = 0;
#pragma unroll
  for (int l = 0; l < 5; ++l) {
    pos = frame_apply(c0, c1, c2, pos, M.origin[l][0], M.origin[l][1], M.origin[l][2]);
    const float a = l < D ? th[l < D ? l : 0] : 0.0f; // no wrist_rotate angle with d = 4 (D3)
    float s, c;
    if (FAST)
      dm::sincos_small(a, &s, &c);
    else
      dm::sincos(a, &s, &c);
    const int ax = M.axis[l];
    if (ax == 0)
      frame_rotate(s, c, c1, c2);
    else if (ax == 1)
      frame_rotate(s, c, c2, c0);
    else
      frame_rotate(s, c, c0, c1);
    const uint32_t qe = M.first[l + 1];
    for (; q < qe; ++q)
      visit(q, frame_apply(c0, c1, c2, pos, M.centre[q][0], M.centre[q][1], M.centre[q][2]));
  }
}
template <int D> struct WpX {
  float x[D];
};
template <int D, int GM, bool FIDX, bool FAST>
__device__ __forceinline__ int model_collisions(const CostParams &P, WpX<D> in) {
  int cnt = 0;
  const FkModelDev &M = P.fk;
  fk_model_walk<D, FAST>(M, in.x, [&](uint32_t q, P3 sp) {
    cnt += collides<GM, FIDX>(P, sp, M.thr[q]);
  });
  return cnt;
}

// One waypoint of one sample: derivatives, smoothness / limit terms and the collision count
// (generate_trajectories.cu:86-103,276-291,391-466).  FAST = straight-line code: reciprocal
// divides and small-argument sin/cos computed unconditionally.  The return value says whether
// the waypoint's joint angles are "fine" (dm::fine_angle): the FAST results are valid when
// this and the two previous waypoints are (device_math.cuh), otherwise the caller redoes the
// waypoint with FAST = false (IEEE divides, full-range sin/cos).  Both produce the same bits
// wherever FAST is valid.  vnew = this waypoint's velocities (the caller commits them).
template <int D, int GM, bool FIDX, bool LIM, bool FAST, bool FKM = false>
__device__ __forceinline__ bool waypoint_cost(const CostParams &P, const float (&x)[D],
                                              const float (&xprev)[D], const float (&vprev)[D],
                                              float (&vnew)[D], float *a2_out, float *l2_out,
                                              int *coll_out) {
  bool fine = true;
  float a2 = 0.0f, l2 = 0.0f;
  auto dv = [&](float a) -> float {
    if (FAST) return dm::div_fast_unguarded(a, P.dt_div);
    return dm::div_(a, P.dt_div.b);
  };
#pragma unroll
  for (int j = 0; j < D; ++j) {
    if (FAST) fine = fine & dm::fine_angle(x[j]);
    const float v = dv(dm::sub_(x[j], xprev[j])); // :288
    const float a = dv(dm::sub_(v, vprev[j]));
    a2 = dm::add_(a2, dm::mul_(a, a)); // :76
    vnew[j] = v;
    if (LIM) {
      const float h =
          fmaxf(0.0f, fmaxf(dm::sub_(x[j], P.lim_max[j]), dm::sub_(P.lim_min[j], x[j])));
      l2 = dm::add_(l2, dm::mul_(h, h));
    }
  }
  int cnt = 0;
  if (GM != GM_NONE) {
    if (FKM) { // FK model kernels only: the reference-chain kernels keep their register budget
      WpX<D> in;
#pragma unroll
      for (int j = 0; j < D; ++j) in.x[j] = x[j];
      cnt = model_collisions<D, GM, FIDX && FAST, FAST>(P, in);
    } else {
      P3 pts[5];
      fk_points<D, FAST>(x, P.rpy_s, P.rpy_c, pts);
#pragma unroll
      for (int l = 0; l < 5; ++l) cnt += collides<GM, FIDX && FAST>(P, pts[l], P.coll_dist);
    }
  }
  *a2_out = a2;
  *l2_out = l2;
  *coll_out = cnt;
  return fine;
}

// by-value interface of the out-of-line general path: keeps the hot loop's arrays in registers
template <int D> struct WpIn {
  float x[D], xprev[D], vprev[D];
};
template <int D> struct WpOut {
  float vnew[D], a2, l2;
  int coll;
};
template <int D, int GM, bool FIDX, bool LIM, bool FKM = false>
__device__ __noinline__ WpOut<D> waypoint_cost_general(const CostParams &P, WpIn<D> in) {
  WpOut<D> o;
  waypoint_cost<D, GM, FIDX, LIM, false, FKM>(P, in.x, in.xprev, in.vprev, o.vnew, &o.a2, &o.l2,
                                         &o.coll);
  return o;
}

__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    unsigned long long o = __shfl_xor_sync(kFull, v, off);
    v = o < v ? o : v;
  }
  return v;
}

// ---- pending update of the PREVIOUS launch, applied while the trajectory is staged -------
// Keep-best (generate_trajectories.cu:225-241, D5/D6 fixed): the previous cost launch left the
// climber's winning key; instead of a separate update kernel every CTA adds the winner's noise
// while it stages the trajectory, and the climber's first CTA also writes the updated state for
// the launch after this one.  Weighted (:117-142): the update vector is subtracted instead.
struct Pending {
  bool improve;  // block-uniform: the staged trajectory gets swin[] / the winner's noise added
  bool weighted;
  uint32_t tg;      // winner's global sample index
  const float *pnz; // winner's noise row in the previous launch's external buffer
};

template <int D>
__device__ __forceinline__ Pending pending_decide(const CostParams &P, uint32_t c, float *swin) {
  Pending pd{false, P.pend_mode == PEND_WEIGHTED, 0u, nullptr};
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (pd.weighted) {
    pd.improve = true;
    if (blockIdx.x == 0 && threadIdx.x == 0) P.found_better[c] = 1;
  } else if (P.pend_mode != PEND_NONE) {
    const unsigned long long pk = __ldg(P.pend_keys + c);
    const uint32_t hi = (uint32_t)(pk >> 32);
    pd.tg = (uint32_t)pk;
    uint32_t cur_new = hi;
    if (P.pend_mode == PEND_KEEP_BEST) {
      const uint32_t cur = __ldg(P.cur_hi_in + c);
      pd.improve = hi < cur;
      cur_new = pd.improve ? hi : cur;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      P.cur_hi_out[c] = cur_new;
      if (P.pend_mode == PEND_SEED)
        P.found_better[c] = 0;
      else if (pd.improve)
        P.found_better[c] = 1;
    }
  }
  if (blockIdx.x == 0 && P.usum_reset && threadIdx.x < 8) P.usum_reset[(size_t)c * 8 + threadIdx.x] = 0;
  if (blockIdx.x == 0 && P.keys_reset && threadIdx.x == 0) P.keys_reset[c] = kKeyNone;
  if (pd.improve && !pd.weighted && P.pend_noise_mode == NOISE_EXTERNAL)
    pd.pnz = P.pend_noise + ((size_t)c * P.noise_m + (size_t)(pd.tg - P.noise_t0)) * P.n_noise * D;
  if (pd.weighted) {
    // trajectory[w][j] -= gain * u[j] (generate_trajectories.cu:139-140): swin[j] = -(gain*u[j])
    if (warp == (blockDim.x >> 5) - 1 && lane < D) {
      const long long sum = P.pend_usum[(size_t)c * 8 + lane];
      const long long poisoned = P.pend_usum[(size_t)c * 8 + 7];
      const float u = poisoned ? __int_as_float(0x7fc00000)
                               : __double2float_rn(__dmul_rn(__ll2double_rn(sum), 0x1p-28));
      swin[lane] = -dm::mul_(P.update_gain, u);
    }
  } else if (pd.improve && P.n_noise == 1 && warp == (blockDim.x >> 5) - 1 && lane < 3) {
    // the winner's noise vector (one d-vector per sample, D16): three lanes of the CTA's last
    // warp redraw its Box-Muller pairs
    float a = 0.0f, b = 0.0f;
    if (P.pend_noise_mode == NOISE_PHILOX) {
      dm::noise_pair(P.seed_lo, P.seed_hi, P.sigma, P.pend_iteration, c, pd.tg, 0, lane, &a, &b);
    } else if (P.pend_noise_mode == NOISE_EXTERNAL) {
      a = 2 * lane < D ? __ldg(pd.pnz + 2 * lane) : 0.0f;
      b = 2 * lane + 1 < D ? __ldg(pd.pnz + 2 * lane + 1) : 0.0f;
    }
    if (2 * lane < D) swin[2 * lane] = a;
    if (2 * lane + 1 < D) swin[2 * lane + 1] = b;
  }
  return pd;
}

// Stage the climber's trajectory (with the pending update applied): SoA [j][npad] in global ->
// AoS [w][DP] in shared; SOA stores [j][npad + 2] instead (waypoint n duplicates n - 1), so
// that one 64-bit shared load yields the packed (w, w + 1) operand of the packed kernel.
template <int D, int DP, bool SOA>
__device__ __forceinline__ void stage_trajectory(const CostParams &P, uint32_t c, const Pending &pd,
                                                 const float *swin, float *sbase) {
  if (pd.improve) __syncthreads(); // block-uniform; swin is complete
  const float *traj = P.traj + (size_t)c * D * P.npad;
  float *tnext = (blockIdx.x == 0 && P.traj_next) ? P.traj_next + (size_t)c * D * P.npad : nullptr;
  for (uint32_t i = threadIdx.x; i < P.n * D; i += blockDim.x) {
    const uint32_t j = i / P.n, w = i - j * P.n;
    float b = __ldg(traj + (size_t)j * P.npad + w);
    if (pd.improve) {
      float e;
      if (P.n_noise == 1)
        e = swin[j];
      else if (P.pend_noise_mode == NOISE_EXTERNAL)
        e = __ldg(pd.pnz + (size_t)w * D + j);
      else
        e = dm::noise_at(P.seed_lo, P.seed_hi, P.sigma, P.pend_iteration, c, pd.tg, w, j);
      b = dm::add_(b, e);
    }
    if (SOA) {
      sbase[j * (P.npad + 2) + w] = b;
      if (w == P.n - 1) sbase[j * (P.npad + 2) + w + 1] = b;
    } else {
      sbase[w * DP + j] = b;
    }
    if (tnext) tnext[(size_t)j * P.npad + w] = b;
  }
  __syncthreads();
}

template <int D, int GM, bool NPW, bool FIDX, bool LIM, bool FKM>
__global__ void __launch_bounds__(kMaxCostThreads)
    cost_kernel(const __grid_constant__ CostParams P) {
  namespace cg = cooperative_groups;
  constexpr int DP = D == 4 ? 4 : 8; // padded joints per waypoint in shared memory
  extern __shared__ __align__(16) float smem[];
  const uint32_t T = 1u << P.parts_log2;   // threads (parts) per sample
  const uint32_t PW = 1u << P.pw_log2;     // parts per CTA; cluster size = T / PW
  const uint32_t cs_log2 = P.parts_log2 - P.pw_log2;
  const uint32_t spb = (blockDim.x >> 5 >> P.pw_log2) * 32; // samples per block (32 if T > 1)
  float *sbase = smem;                                 // [n][DP]
  float *seps = sbase + (size_t)P.n * DP;              // [D][32]      (shared noise, PW >= 4)
  float *swin = seps + (PW >= 4 ? D * 32 : 0);         // [8]          pending winner's noise
  float *spart = swin + 8;                             // [n][32]      (T > 1; rank-0 CTA's is used)
  float *slim = spart + (T > 1 ? (size_t)P.n * 32 : 0); // [n][32]     (T > 1 and limits)
  int *scoll = (int *)(slim + ((T > 1 && LIM) ? (size_t)P.n * 32 : 0)); // [T][32]

  const bool clustered = cs_log2 != 0;
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t crank = clustered ? cluster.block_rank() : 0u;
  if (clustered) cluster.barrier_arrive(); // every CTA of the cluster is running before DSMEM use

  const uint32_t c = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t part = T > 1 ? crank * PW + warp : 0u;
  const uint32_t si = T > 1 ? (uint32_t)lane : warp * 32 + lane; // sample slot inside the block
  uint32_t t = (blockIdx.x >> cs_log2) * spb + si;
  const bool active = t < P.m;
  t = active ? t : P.m - 1; // idle lanes shadow the last sample (no divergence, no stores)
  const uint32_t tg = (uint32_t)(P.sample_offset + t);

  float start[D];
#pragma unroll
  for (int j = 0; j < D; ++j) start[j] = __ldg(P.start + j);

  const Pending pd = pending_decide<D>(P, c, swin);

  const float *nz = nullptr;
  if (P.noise_mode == NOISE_EXTERNAL)
    nz = P.noise + ((size_t)c * P.noise_m + (size_t)(tg - P.noise_t0)) * P.n_noise * D;
  float eps[D];
#pragma unroll
  for (int j = 0; j < D; ++j) eps[j] = 0.0f;
  const bool share_noise = !NPW && PW >= 4 && P.noise_mode == NOISE_PHILOX;
  if (!NPW) {
    if (P.noise_mode == NOISE_EXTERNAL) {
#pragma unroll
      for (int j = 0; j < D; ++j) eps[j] = __ldg(nz + j);
    } else if (P.noise_mode == NOISE_PHILOX) {
      if (!share_noise) {
        dm::noise_vec<D>(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, 0, eps);
      } else if (warp < 3) {
        // the CTA's warps all serve the same 32 samples: warps 0/1/2 draw the Box-Muller
        // pairs (joints 0,1), (2,3) and joint 4 once and share them through shared memory
        float a, b;
        dm::noise_pair(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, 0, warp, &a, &b);
        if (2 * warp < D) seps[(2 * warp) * 32 + lane] = a;
        if (2 * warp + 1 < D && warp < 2) seps[(2 * warp + 1) * 32 + lane] = b;
      }
    }
  }
  stage_trajectory<D, DP, false>(P, c, pd, swin, sbase);
  if (share_noise) {
#pragma unroll
    for (int j = 0; j < D; ++j) eps[j] = seps[j * 32 + lane];
  }

  // noisy position of joint j at waypoint w (generate_trajectories.cu:53-57)
  auto noisy = [&](uint32_t w, int j) -> float {
    float e = eps[j];
    if (NPW) {
      if (P.noise_mode == NOISE_EXTERNAL)
        e = __ldg(nz + (size_t)w * D + j);
      else if (P.noise_mode == NOISE_PHILOX)
        e = dm::noise_at(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, w, j);
    }
    return dm::add_(sbase[w * DP + j], e);
  };

  uint32_t w = part * P.wpp;
  uint32_t w_end = w + P.wpp;
  w_end = w_end < P.n ? w_end : P.n;

  // priors of the first owned waypoint: the initial waypoint for BOTH derivative
  // passes at w = 0 (generate_trajectories.cu:285), else recomputed.
  // fine1 / fine2: were the angles of the previous two waypoints "fine" (dm::fine_angle)?
  // The straight-line results of a waypoint are used only if it and both predecessors are.
  float xprev[D], vprev[D];
  bool fine1 = P.fast_ok != 0, fine2 = fine1;
  if (w == 0) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      xprev[j] = vprev[j] = start[j];
      fine1 = fine1 & dm::fine_angle(start[j]);
    }
  } else if (w < P.n) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const float x1 = noisy(w - 1, j);
      const float x2 = w >= 2 ? noisy(w - 2, j) : start[j];
      xprev[j] = x1;
      vprev[j] = dm::div_fast(dm::sub_(x1, x2), P.dt_div);
      fine1 = fine1 & dm::fine_angle(x1);
      fine2 = fine2 & dm::fine_angle(x2);
    }
  }

  // partial sums of the other CTAs of the cluster land in rank 0's shared memory (DSMEM)
  float *spart0 = spart, *slim0 = slim;
  int *scoll0 = scoll;
  if (clustered) {
    spart0 = cluster.map_shared_rank(spart, 0);
    slim0 = cluster.map_shared_rank(slim, 0);
    scoll0 = cluster.map_shared_rank(scoll, 0);
    cluster.barrier_wait();
  }

  float total = 0.0f, ltotal = 0.0f;
  int coll = 0;
#pragma unroll 2
  for (; w < w_end; ++w) {
    float x[D], vnew[D];
#pragma unroll
    for (int j = 0; j < D; ++j) x[j] = noisy(w, j);
    float a2, l2;
    int cw;
    const bool fine0 = waypoint_cost<D, GM, FIDX, LIM, true, FKM>(P, x, xprev, vprev, vnew, &a2, &l2, &cw);
    const bool fast_valid = fine0 & fine1 & fine2;
    fine2 = fine1;
    fine1 = fine0 & (P.fast_ok != 0);
    if (!fast_valid) {
      WpIn<D> in; // rare: an angle outside the verified ranges -> IEEE divides, full sin/cos
#pragma unroll
      for (int j = 0; j < D; ++j) {
        in.x[j] = x[j];
        in.xprev[j] = xprev[j];
        in.vprev[j] = vprev[j];
      }
      const WpOut<D> o = waypoint_cost_general<D, GM, FIDX, LIM, FKM>(P, in);
#pragma unroll
      for (int j = 0; j < D; ++j) vnew[j] = o.vnew[j];
      a2 = o.a2;
      l2 = o.l2;
      cw = o.coll;
    }
#pragma unroll
    for (int j = 0; j < D; ++j) {
      xprev[j] = x[j];
      vprev[j] = vnew[j];
    }
    coll += cw;
    if (T == 1) { // fixed order: left fold over the waypoints (oracle ORC_SUM_WAYPOINT)
      total = dm::add_(total, a2);
      if (LIM) ltotal = dm::add_(ltotal, l2);
    } else {
      spart0[w * 32 + lane] = a2;
      if (LIM) slim0[w * 32 + lane] = l2;
    }
  }
  if (T > 1) {
    scoll0[part * 32 + lane] = coll;
    if (clustered)
      cluster.sync();
    else
      __syncthreads();
    if (part != 0) return;
    coll = 0;
    for (uint32_t q = 0; q < T; ++q) coll += scoll[q * 32 + lane];
    for (uint32_t ww = 0; ww < P.n; ++ww) {
      total = dm::add_(total, spart[ww * 32 + lane]);
      if (LIM) ltotal = dm::add_(ltotal, slim[ww * 32 + lane]);
    }
  }

  float score = dm::add_((float)coll, dm::mul_(P.smooth_scale, total)); // :81,100,107-108
  if (LIM && P.limit_weight != 0.0f) score = dm::add_(score, dm::mul_(P.limit_weight, ltotal));
  if (active) {
    const size_t s = (size_t)c * P.m + t;
    if (P.scores) P.scores[s] = score;
    if (P.collisions) P.collisions[s] = coll;
  }
  if (P.keys_float) {
    unsigned long long key =
        active ? ((unsigned long long)key_hi_float(score) << 32) | tg : kKeyNone;
    key = warp_min_u64(key);
    if (lane == 0) atomicMin(P.keys_float + c, key);
  }
  if (P.keys) {
    unsigned long long key =
        active ? ((unsigned long long)key_hi(P.select_mode, score) << 32) | tg : kKeyNone;
    key = warp_min_u64(key);
    if (lane == 0) atomicMin(P.keys + c, key);
  }
}

// ================================================================ persistent planner kernel
// Small problems (BASELINE config 2: 1024 samples x 64 waypoints) are latency-bound: a launch
// per iteration costs more than the iteration's arithmetic.  Like the reference's
// optimize_trajectories (generate_trajectories.cu:197-253: one launch, loop inside), this kernel
// runs the whole keep-best loop in one cooperative launch: a CTA owns a few samples with
// 2^t_log2 threads each (one or a few waypoints per thread), the trajectory stays in shared
// memory across iterations, and one grid-wide barrier per iteration publishes the argmin key.
// Arithmetic, noise, keys and tie-breaking are those of cost_kernel, so the result is
// bit-identical to the launch-per-iteration path.
// Per-iteration, per-climber barrier word: 16 bytes {argmin key, arrivals - 1}, all ones when
// armed.  A CTA publishes with atomicMin on the key, a fence, then atomicAdd on the count; the
// waiter polls BOTH with one 16-byte load (one transaction on one L2 sector): when the count is
// complete every atomicMin ordered before an arrival has been performed on that sector, so the
// key half of the same load is the final minimum -- no second round trip to re-read the key.
// The key is the only data that crosses CTAs here, so no acquire fence follows the poll.
__device__ __forceinline__ void grid_publish(unsigned long long *word, unsigned long long key) {
  atomicMin(word, key);
  __threadfence();
  atomicAdd(word + 1, 1ull);
}
__device__ __forceinline__ unsigned long long grid_wait_key(const unsigned long long *word,
                                                            unsigned long long arrivals) {
  unsigned long long key, cnt;
  do {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(key), "=l"(cnt) : "l"(word) : "memory");
  } while (cnt + 1 < arrivals); // armed: cnt + 1 == 0
  return key;
}

template <int D, int GM, bool LIM, bool FKM>
__global__ void __launch_bounds__(kPersistThreads) plan_persistent_kernel(const __grid_constant__ PersistParams Q) {
  const CostParams &P = Q.cost;
  constexpr bool FIDX = GM == GM_U8_CDG || GM == GM_F32_CDG;
  extern __shared__ __align__(16) float smem[];
  const uint32_t T = 1u << Q.t_log2, SB = kPersistComputeThreads >> Q.t_log2;
  const uint32_t WS = T * P.wpp; // waypoint slots per sample (>= n)
  // warps 0..7 cost the samples; warp 8 (the helper) does the grid-wide key exchange meanwhile
  const bool helper = threadIdx.x >= kPersistComputeThreads;
  const uint32_t hl = threadIdx.x - kPersistComputeThreads; // helper lane
  float *sbase = smem;                                   // [D][npad]  current trajectory
  float *spart = sbase + (size_t)D * P.npad;             // [SB][WS]   per-waypoint sum of acc^2
  float *slim = spart + (size_t)SB * WS;                 // [SB][WS]   (limits)
  float *seps = slim + (LIM ? (size_t)SB * WS : 0);      // [2][SB][8] noise of iteration it in half it & 1
  float *swin = seps + (size_t)SB * 16;                  // [8]        winner's noise
  int *scoll = (int *)(swin + 8);                        // [SB]
  unsigned long long *skey = (unsigned long long *)(scoll + ((SB + 1) & ~1u)); // [2] CTA minimum of iteration it in slot it & 1
  unsigned long long *sbest = skey + 2;                                        // [2] merged minimum of iteration it in slot it & 1

  const uint32_t c = blockIdx.y, k = gridDim.y;
  const uint32_t s = threadIdx.x >> Q.t_log2, r = threadIdx.x & (T - 1);
  uint32_t t = blockIdx.x * SB + s;
  const bool active = t < P.m;
  t = active ? t : P.m - 1;
  const uint32_t tg = (uint32_t)(P.sample_offset + t);

  float start[D];
#pragma unroll
  for (int j = 0; j < D; ++j) start[j] = __ldg(P.start + j);
  for (uint32_t i = threadIdx.x; i < (uint32_t)D * P.npad; i += blockDim.x)
    sbase[i] = __ldg(P.traj + (size_t)c * D * P.npad + i);
  if (threadIdx.x < SB) scoll[threadIdx.x] = 0;
  if (threadIdx.x < 2) skey[threadIdx.x] = kKeyNone;
  // the seed pass scored the unperturbed trajectory (generate_trajectories.cu:180-191)
  uint32_t cur_hi = (uint32_t)(__ldg(P.pend_keys + c) >> 32);
  bool found = false;

  // one d-vector of noise per sample and iteration (generate_trajectories.cu:37-44): the
  // sample's first three threads draw one Box-Muller pair each
  // (threads T/2 .. T/2 + 2 of the sample when T >= 8: with T = 64 a different warp than the one whose
  // lane 0 folds the sample's sums, so the two overlap)
  const uint32_t rn = r - (T >= 8 ? (T >> 1) : 0u); // 0, 1, 2 on the noise threads
  auto gen_noise = [&](uint32_t it) {
    if (helper) return;
    float *se = seps + (size_t)(it & 1) * SB * 8 + s * 8;
    if (P.noise_mode == NOISE_PHILOX) {
      if (T >= 4) {
        if (rn < 3) {
          float a, b;
          dm::noise_pair(P.seed_lo, P.seed_hi, P.sigma, it, c, tg, 0, rn, &a, &b);
          if (2 * rn < D) se[2 * rn] = a;
          if (2 * rn + 1 < D) se[2 * rn + 1] = b;
        }
      } else if (r == 0) {
        float e[D];
        dm::noise_vec<D>(P.seed_lo, P.seed_hi, P.sigma, it, c, tg, 0, e);
#pragma unroll
        for (int j = 0; j < D; ++j) se[j] = e[j];
      }
    } else if (r == 0) {
#pragma unroll
      for (int j = 0; j < D; ++j)
        se[j] =
            P.noise_mode == NOISE_EXTERNAL
                ? __ldg(Q.noise_base + (size_t)it * Q.noise_iter_stride +
                        ((size_t)c * P.noise_m + (size_t)(tg - P.noise_t0)) * D + j)
                : 0.0f;
    }
  };
  // the mailbox set of the NEXT launch: nobody writes it before this rank's last key of this
  // launch has left
  if (Q.nranks > 1 && blockIdx.x == 0 && threadIdx.x < 24)
    Q.mailbox_next[((size_t)(threadIdx.x >> 3) * k + c) * 8 + (threadIdx.x & 7)] = kKeyNone;
  gen_noise(0);
  __syncthreads();

  const uint32_t w0 = r * P.wpp;
  const uint32_t w1 = (w0 + P.wpp) < P.n ? (w0 + P.wpp) : P.n;

  // Helper warp: sends the CTA's key of iteration it (taken from *skey into h_mine earlier)
  // into the grid-wide argmin, waits for the merged key -- with several ranks through the
  // NVLink peer mailboxes -- and leaves it in sbest[it & 1].
  unsigned long long h_mine = kKeyNone;
  auto helper_exchange = [&](uint32_t it) {
    unsigned long long *kcur = Q.keys3 + ((size_t)(it % 3) * k + c) * 2;
    if (hl == 0) {
      // re-arm the word of iteration it + 1: its last readers passed barrier it - 1
      if (blockIdx.x == 0)
        *reinterpret_cast<ulonglong2 *>(Q.keys3 + ((size_t)((it + 1) % 3) * k + c) * 2) =
            make_ulonglong2(kKeyNone, kKeyNone);
      grid_publish(kcur, h_mine);
    }
    if (Q.nranks <= 1) {
      if (hl == 0) sbest[it & 1] = grid_wait_key(kcur, gridDim.x);
    } else {
      // The climber's first CTA waits for the rank's CTAs, then stores the rank's key into the
      // mailbox of every rank (its own included): one 8-byte store per peer over NVLink,
      // payload = flag.  EVERY CTA polls the rank's own mailbox -- local memory -- for the
      // nranks keys of this iteration and takes the minimum.  Buffers rotate three ways; the
      // slots of iteration it + 1 are re-armed before this rank's key of iteration it leaves,
      // so no peer can be writing them yet and no local CTA is still reading them.
      const size_t cur = ((size_t)(it % 3) * k + c) * 8;
      if (blockIdx.x == 0) {
        unsigned long long local = 0;
        if (hl == 0) local = grid_wait_key(kcur, gridDim.x);
        local = __shfl_sync(kFull, local, 0);
        if (hl < Q.nranks) {
          Q.mailbox[Q.rank][((size_t)((it + 1) % 3) * k + c) * 8 + hl] = kKeyNone;
          asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(Q.mailbox[hl] + cur + Q.rank), "l"(local) : "memory");
        }
      }
      unsigned long long v = kKeyNone;
      if (hl < Q.nranks) {
        const unsigned long long *mine = Q.mailbox[Q.rank] + cur + hl;
        do {
          asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
        } while (v == kKeyNone);
      }
      v = warp_min_u64(v);
      if (hl == 0) sbest[it & 1] = v;
    }
  };

  // One iteration's cost pass over the CTA's samples with the noise in seps and the trajectory
  // in sbase; leaves the CTA's minimum key in *skey.  While the eight compute warps do that, the
  // helper warp runs the key exchange of the PREVIOUS iteration (exchange_prev).
  auto evaluate = [&](uint32_t ev, bool exchange_prev) {
    if (helper) {
      if (exchange_prev) helper_exchange(ev - 1);
    } else {
      float eps[D];
#pragma unroll
      for (int j = 0; j < D; ++j) eps[j] = seps[(size_t)(ev & 1) * SB * 8 + s * 8 + j];
      auto noisy = [&](uint32_t w, int j) -> float { return dm::add_(sbase[j * P.npad + w], eps[j]); };
      uint32_t w = w0;
      float xprev[D], vprev[D];
      bool fine1 = P.fast_ok != 0, fine2 = fine1;
      if (w == 0) {
#pragma unroll
        for (int j = 0; j < D; ++j) {
          xprev[j] = vprev[j] = start[j];
          fine1 = fine1 & dm::fine_angle(start[j]);
        }
      } else if (w < P.n) {
#pragma unroll
        for (int j = 0; j < D; ++j) {
          const float x1 = noisy(w - 1, j);
          const float x2 = w >= 2 ? noisy(w - 2, j) : start[j];
          xprev[j] = x1;
          vprev[j] = dm::div_fast(dm::sub_(x1, x2), P.dt_div);
          fine1 = fine1 & dm::fine_angle(x1);
          fine2 = fine2 & dm::fine_angle(x2);
        }
      }
      int coll = 0;
      for (; w < w1; ++w) {
        float x[D], vnew[D];
#pragma unroll
        for (int j = 0; j < D; ++j) x[j] = noisy(w, j);
        float a2, l2;
        int cw;
        const bool fine0 = waypoint_cost<D, GM, FIDX, LIM, true, FKM>(P, x, xprev, vprev, vnew, &a2, &l2, &cw);
        const bool fast_valid = fine0 & fine1 & fine2;
        fine2 = fine1;
        fine1 = fine0 & (P.fast_ok != 0);
        if (!fast_valid) {
          WpIn<D> in;
#pragma unroll
          for (int j = 0; j < D; ++j) {
            in.x[j] = x[j];
            in.xprev[j] = xprev[j];
            in.vprev[j] = vprev[j];
          }
          const WpOut<D> o = waypoint_cost_general<D, GM, FIDX, LIM, FKM>(P, in);
#pragma unroll
          for (int j = 0; j < D; ++j) vnew[j] = o.vnew[j];
          a2 = o.a2;
          l2 = o.l2;
          cw = o.coll;
        }
#pragma unroll
        for (int j = 0; j < D; ++j) {
          xprev[j] = x[j];
          vprev[j] = vnew[j];
        }
        coll += cw;
        spart[s * WS + w] = a2;
        if (LIM) slim[s * WS + w] = l2;
      }
      if (coll) atomicAdd(scoll + s, coll); // integer: order-independent
      // the compute warps only: the helper joins at the end of the pass, so its exchange also
      // overlaps the fold
      asm volatile("bar.sync 1, %0;" ::"n"(kPersistComputeThreads) : "memory");
      if (ev + 1 < Q.iters) gen_noise(ev + 1); // into the other half of seps
    }
    if (!helper && r == 0) { // fixed order: left fold over the waypoints (oracle ORC_SUM_WAYPOINT)
      float total = 0.0f, ltotal = 0.0f;
      for (uint32_t ww = 0; ww < P.n; ++ww) {
        total = dm::add_(total, spart[s * WS + ww]);
        if (LIM) ltotal = dm::add_(ltotal, slim[s * WS + ww]);
      }
      const int ctot = scoll[s];
      scoll[s] = 0;
      float score = dm::add_((float)ctot, dm::mul_(P.smooth_scale, total));
      if (LIM && P.limit_weight != 0.0f) score = dm::add_(score, dm::mul_(P.limit_weight, ltotal));
      if (active)
        atomicMin(skey + (ev & 1), ((unsigned long long)key_hi(P.select_mode, score) << 32) | tg);
    }
    __syncthreads();
  };

  // Applies the keep-best of iteration it (merged key in *sbest) to the CTA's copy of the
  // trajectory (generate_trajectories.cu:225-241, D5/D6 fixed; merged key in sbest[it & 1]).  Returns whether the trajectory
  // changed (block-uniform).
  auto apply = [&](uint32_t it) -> bool {
    const unsigned long long best = sbest[it & 1];
    const uint32_t hi = (uint32_t)(best >> 32), win = (uint32_t)best;
    if (hi >= cur_hi) return false;
    cur_hi = hi;
    found = true;
    if (threadIdx.x < 3) {
      float a = 0.0f, b = 0.0f;
      const uint32_t q = threadIdx.x;
      if (P.noise_mode == NOISE_PHILOX) {
        dm::noise_pair(P.seed_lo, P.seed_hi, P.sigma, it, c, win, 0, q, &a, &b);
      } else if (P.noise_mode == NOISE_EXTERNAL) {
        const float *pn = Q.noise_base + (size_t)it * Q.noise_iter_stride +
                          ((size_t)c * P.noise_m + (size_t)(win - P.noise_t0)) * D;
        a = 2 * q < D ? __ldg(pn + 2 * q) : 0.0f;
        b = 2 * q + 1 < D ? __ldg(pn + 2 * q + 1) : 0.0f;
      }
      if (2 * q < D) swin[2 * q] = a;
      if (2 * q + 1 < D) swin[2 * q + 1] = b;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < (uint32_t)D * P.npad; i += blockDim.x) {
      const uint32_t j = i / P.npad;
      if (i - j * P.npad < P.n) sbase[i] = dm::add_(sbase[i], swin[j]);
    }
    __syncthreads();
    return true;
  };

  // Software-pipelined keep-best loop.  Most iterations do not improve the trajectory, so the
  // cost pass of iteration ev runs SPECULATIVELY on the current trajectory while the helper
  // warp exchanges the keys of iteration ev - 1 (two L2 round trips, or NVLink ones, that the
  // compute warps never wait for).  If that exchange does bring an update, the pass is repeated
  // on the updated trajectory (same noise) before the key of iteration ev is taken.  The keys
  // that leave the CTA are therefore exactly those of the sequential loop.
  // No CTA-wide barrier separates two passes: noise and CTA key are double-buffered by the
  // iteration's parity, and the helper re-arms a key slot before the barrier that ends the
  // next pass, i.e. before the slot's next fold.
  if (Q.iters > 0) {
    uint32_t ev = 0;
    bool pending = false; // the key of iteration ev - 1 is taken but not yet exchanged
    for (;;) {
      evaluate(ev, pending);
      if (pending) {
        pending = false;
        if (apply(ev - 1)) { // mis-speculated: discard, evaluate ev again
          if (threadIdx.x == 0) skey[ev & 1] = kKeyNone;
          __syncthreads();
          continue;
        }
      }
      if (helper && hl == 0) { // the key of iteration ev is final: take it
        h_mine = skey[ev & 1];
        skey[ev & 1] = kKeyNone;
      }
      pending = true;
      if (ev + 1 == Q.iters) {
        if (helper) helper_exchange(ev);
        __syncthreads();
        apply(ev);
        break;
      }
      ++ev;
    }
  }
  if (blockIdx.x == 0) { // the climber's final state, for the final scoring pass
    for (uint32_t i = threadIdx.x; i < (uint32_t)D * P.npad; i += blockDim.x)
      P.traj_next[(size_t)c * D * P.npad + i] = sbase[i];
    if (threadIdx.x == 0) {
      P.cur_hi_out[c] = cur_hi;
      P.found_better[c] = found ? 1 : 0;
    }
  }
}

// ===================================================== packed (two samples per thread) path
// For launches with one thread per sample, the thread walks its waypoints two at a time and
// all fp32 arithmetic is issued as packed fma.rn.f32x2 (FFMA2 / FADD2 / FMUL2), waypoint w in
// the low half and w + 1 in the high half:
// the FMA pipe does the same work in half the issue slots, which is what bounds this kernel
// (profiles/r01).  Per-half work is only what has no packed form: compares, float->int,
// index arithmetic, the gathers.  Results equal the scalar kernel bit for bit.
struct P3v {
  dm::f2 x, y, z;
};
__device__ __forceinline__ P3v rot_y2(dm::f2 s, dm::f2 c, P3v q) {
  P3v u;
  u.x = dm::f2_fma(s, q.z, dm::f2_mul(c, q.x));
  u.y = q.y;
  u.z = dm::f2_fma(c, q.z, dm::f2_mul(dm::f2_neg(s), q.x));
  return u;
}
__device__ __forceinline__ P3v rot_z2(dm::f2 s, dm::f2 c, P3v q) {
  P3v u;
  u.x = dm::f2_fma(dm::f2_neg(s), q.y, dm::f2_mul(c, q.x));
  u.y = dm::f2_fma(c, q.y, dm::f2_mul(s, q.x));
  u.z = q.z;
  return u;
}

// fk_points<D, true> on two waypoints (generate_trajectories.cu:397-463)
template <int D>
__device__ __forceinline__ void fk_points2(const dm::f2 (&th)[D], float rpy_s, float rpy_c,
                                           P3v (&out)[5]) {
  using namespace dm;
  f2 s, c;
  P3v p, q;
  f2_sincos_small(th[0], &s, &c); // base, gt.cu:405-410
  q = P3v{f2_splat(0.0f), f2_splat(-0.0015f), f2_splat(0.0715f)};
  p = rot_y2(s, c, q);
  out[0] = p;
  f2_sincos_small(th[1], &s, &c); // shoulder, :416-421
  q = P3v{p.x, p.y, f2_add(p.z, f2_splat(0.045f))};
  p = rot_z2(s, c, q);
  out[1] = p;
  f2_sincos_small(th[2], &s, &c); // elbow, :427-432
  q = P3v{p.x, p.y, f2_add(p.z, f2_splat(0.1f))};
  p = rot_y2(s, c, q);
  out[2] = p;
  f2_sincos_small(th[3], &s, &c); // wrist, :438-443
  q = P3v{p.x, p.y, f2_add(p.z, f2_splat(0.1f))};
  p = rot_y2(s, c, q);
  out[3] = p;
  q = P3v{p.x, f2_add(p.y, f2_splat(-0.0125f)), f2_add(p.z, f2_splat(0.08f))}; // :449-454
  if (D >= 5) {
    f2_sincos_small(th[D >= 5 ? 4 : 0], &s, &c);
    p = rot_y2(s, c, q);
  } else {
    p = q;
  }
  p = rot_y2(f2_splat(rpy_s), f2_splat(rpy_c), p); // :456-459
  p = P3v{p.x, f2_add(p.y, f2_splat(-0.005f)), f2_add(p.z, f2_splat(0.07f))};
  out[4] = p;
}

// collides<GM, true> on two points (threshold form, see GridDev::tx_bits)
template <int GM>
__device__ __forceinline__ void collides2(const CostParams &P, P3v p, float thr, int *hit_lo,
                                          int *hit_hi) {
  using namespace dm;
  const GridDev &g = P.grid;
  const f2 ax = f2_add(p.x, f2_splat(g.x_off)), ay = f2_add(p.y, f2_splat(g.y_off));
  const f2 az = f2_add(p.z, f2_splat(0.0f));
  const f2 fx = f2_div_fast_unguarded(ax, P.res_div), fy = f2_div_fast_unguarded(ay, P.res_div),
           fz = f2_div_fast_unguarded(az, P.res_div);
  auto half = [&](float axh, float ayh, float azh, float fxh, float fyh, float fzh) -> int {
    const bool in = (__float_as_uint(axh) < g.tx_bits) & (__float_as_uint(ayh) < g.ty_bits) &
                    (__float_as_uint(azh) < g.tz_bits);
    const uint32_t ix = __float2uint_rz(fxh), iy = __float2uint_rz(fyh), iz = __float2uint_rz(fzh);
    const uint32_t lin = (iz * g.ny + iy) * g.nx + ix;
    int hit = 0;
    if (GM == GM_U8_CDG) {
      if (in) hit = __ldg((const uint8_t *)g.cells + lin) != 0;
    } else {
      if (in) hit = __ldg((const float *)g.cells + lin) <= thr;
    }
    return hit;
  };
  *hit_lo = half(f2_lo(ax), f2_lo(ay), f2_lo(az), f2_lo(fx), f2_lo(fy), f2_lo(fz));
  *hit_hi = half(f2_hi(ax), f2_hi(ay), f2_hi(az), f2_hi(fx), f2_hi(fy), f2_hi(fz));
}

// fk_model_walk<D, true> + the sphere lookups on two waypoints (frame_apply / frame_rotate in
// packed form: the same multiplies, FMAs and adds per half)
__device__ __forceinline__ P3v frame_apply2(const P3v &c0, const P3v &c1, const P3v &c2,
                                            const P3v &pos, float vx, float vy, float vz) {
  using namespace dm;
  const f2 x2 = f2_splat(vx), y2 = f2_splat(vy), z2 = f2_splat(vz);
  P3v o;
  o.x = f2_add(pos.x, f2_fma(c2.x, z2, f2_fma(c1.x, y2, f2_mul(c0.x, x2))));
  o.y = f2_add(pos.y, f2_fma(c2.y, z2, f2_fma(c1.y, y2, f2_mul(c0.y, x2))));
  o.z = f2_add(pos.z, f2_fma(c2.z, z2, f2_fma(c1.z, y2, f2_mul(c0.z, x2))));
  return o;
}
__device__ __forceinline__ void frame_rotate2(dm::f2 s, dm::f2 c, P3v &b, P3v &cc) {
  using namespace dm;
  const f2 ns = f2_neg(s);
  const P3v nb{f2_fma(s, cc.x, f2_mul(c, b.x)), f2_fma(s, cc.y, f2_mul(c, b.y)),
               f2_fma(s, cc.z, f2_mul(c, b.z))};
  const P3v nc{f2_fma(c, cc.x, f2_mul(ns, b.x)), f2_fma(c, cc.y, f2_mul(ns, b.y)),
               f2_fma(c, cc.z, f2_mul(ns, b.z))};
  b = nb;
  cc = nc;
}
template <int D, int GM>
__device__ __forceinline__ void model_collisions2(const CostParams &P, const dm::f2 (&th)[D],
                                                  int *c_lo, int *c_hi) {
  using namespace dm;
  const FkModelDev &M = P.fk;
  const f2 one = f2_splat(1.0f), zero = f2_splat(0.0f);
  P3v c0{one, zero, zero}, c1{zero, one, zero}, c2{zero, zero, one}, pos{zero, zero, zero};
  int lo = 0, hi = 0;
  uint32_t q = 0;
#pragma unroll
  for (int l = 0; l < 5; ++l) {
    pos = frame_apply2(c0, c1, c2, pos, M.origin[l][0], M.origin[l][1], M.origin[l][2]);
    f2 s, c;
    f2_sincos_small(l < D ? th[l < D ? l : 0] : zero, &s, &c); // no wrist_rotate angle with d = 4
    const int ax = M.axis[l];
    if (ax == 0)
      frame_rotate2(s, c, c1, c2);
    else if (ax == 1)
      frame_rotate2(s, c, c2, c0);
    else
      frame_rotate2(s, c, c0, c1);
    const uint32_t qe = M.first[l + 1];
    for (; q < qe; ++q) {
      int h0, h1;
      collides2<GM>(P, frame_apply2(c0, c1, c2, pos, M.centre[q][0], M.centre[q][1], M.centre[q][2]),
                    M.thr[q], &h0, &h1);
      lo += h0;
      hi += h1;
    }
  }
  *c_lo = lo;
  *c_hi = hi;
}

// waypoint_cost<..., FAST = true> on two CONSECUTIVE WAYPOINTS (lo = w, hi = w + 1) of one
// sample.  xprev / vprev are waypoint w - 1's angles and velocities; the hi half's priors are
// the lo half's own results.  fine_lo / fine_hi as waypoint_cost's return value.
template <int D, int GM, bool LIM, bool FKM>
__device__ __forceinline__ void waypoint_cost2(const CostParams &P, const dm::f2 (&x)[D],
                                               const float (&xprev)[D], const float (&vprev)[D],
                                               dm::f2 (&vnew)[D], dm::f2 *a2_out, dm::f2 *l2_out,
                                               int *coll_lo, int *coll_hi, bool *fine_lo,
                                               bool *fine_hi) {
  using namespace dm;
  bool f_lo = true, f_hi = true;
  f2 a2 = f2_splat(0.0f), l2 = f2_splat(0.0f);
#pragma unroll
  for (int j = 0; j < D; ++j) {
    f_lo = f_lo & fine_angle(f2_lo(x[j]));
    f_hi = f_hi & fine_angle(f2_hi(x[j]));
    const f2 xp = f2_pack(xprev[j], f2_lo(x[j]));
    const f2 v = f2_div_fast_unguarded(f2_sub(x[j], xp), P.dt_div); // :288
    const f2 vp = f2_pack(vprev[j], f2_lo(v));
    const f2 a = f2_div_fast_unguarded(f2_sub(v, vp), P.dt_div);
    a2 = f2_add(a2, f2_mul_rounded(a, a)); // :76 (product rounded, then added)
    vnew[j] = v;
    if (LIM) {
      const f2 over = f2_sub(x[j], f2_splat(P.lim_max[j])), under = f2_sub(f2_splat(P.lim_min[j]), x[j]);
      const float h0 = fmaxf(0.0f, fmaxf(f2_lo(over), f2_lo(under)));
      const float h1 = fmaxf(0.0f, fmaxf(f2_hi(over), f2_hi(under)));
      l2 = f2_add(l2, f2_pack(__fmul_rn(h0, h0), __fmul_rn(h1, h1)));
    }
  }
  int c_lo = 0, c_hi = 0;
  if (GM != GM_NONE && FKM) {
    model_collisions2<D, GM>(P, x, &c_lo, &c_hi);
  } else if (GM != GM_NONE) {
    P3v pts[5];
    fk_points2<D>(x, P.rpy_s, P.rpy_c, pts);
#pragma unroll
    for (int l = 0; l < 5; ++l) {
      int h0, h1;
      collides2<GM>(P, pts[l], P.coll_dist, &h0, &h1);
      c_lo += h0;
      c_hi += h1;
    }
  }
  *a2_out = a2;
  *l2_out = l2;
  *coll_lo = c_lo;
  *coll_hi = c_hi;
  *fine_lo = f_lo;
  *fine_hi = f_hi;
}

constexpr int kPkThreads = 128;
constexpr uint32_t kPkMaxPartN = 256; // T > 1: 32 KB of per-waypoint sums per CTA at most
#ifndef DOAPP_PK_MINBLOCKS
#define DOAPP_PK_MINBLOCKS 4
#endif

// One thread = one sample (or one of 2 / 4 runs of a sample's waypoints, P.parts_log2), TWO
// consecutive waypoints per step in the two halves of packed registers.  Same sample-to-lane
// mapping, noise, staging prologue, sum order and outputs as cost_kernel.  With T = 2 or 4
// parts per sample the 4 warps of a CTA are 4 / T sample-warps x T parts; every part leaves
// its per-waypoint sums in shared memory and part 0 folds them in waypoint order.
template <int D, int GM, bool LIM, bool FKM, bool MP>
__global__ void __launch_bounds__(kPkThreads, DOAPP_PK_MINBLOCKS)
    cost_kernel_pk(const __grid_constant__ CostParams P) {
  using namespace dm;
  extern __shared__ __align__(16) float smem[];
  const uint32_t NP = P.npad + 2;
  float *sbase = smem;                     // [D][NP], waypoint n duplicates n - 1
  float *swin = sbase + (size_t)D * NP;    // [8] pending winner's noise / update vector
  // MP (multi-part) instances: 2 or 4 parts per sample; the one-part instances fold T == 1
  const uint32_t TL = MP ? P.parts_log2 : 0u;
  const uint32_t T = 1u << TL;             // 1, 2 or 4 parts per sample
  const uint32_t WS = T * P.wpp;           // waypoint slots per sample (>= n), T > 1 only
  const uint32_t SW = (kPkThreads >> 5) >> TL; // sample-warps per CTA
  float *spart = swin + 8;                                      // [SW][WS][32]
  float *slim = spart + (T > 1 ? (size_t)SW * WS * 32 : 0);     // [SW][WS][32] (limits)
  int *scoll = (int *)(slim + (T > 1 && LIM ? (size_t)SW * WS * 32 : 0)); // [4][32]
  const uint32_t c = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const uint32_t warp = threadIdx.x >> 5, part = warp & (T - 1), swi = warp >> TL;
  uint32_t t = blockIdx.x * (SW * 32) + swi * 32 + lane;
  const bool active = t < P.m;
  t = active ? t : P.m - 1; // idle lanes shadow the last sample (no divergence, no stores)
  const uint32_t tg = (uint32_t)(P.sample_offset + t);

  const Pending pd = pending_decide<D>(P, c, swin);

  // stage 1: one d-vector of noise per sample (generate_trajectories.cu:37-59)
  float eps[D];
#pragma unroll
  for (int j = 0; j < D; ++j) eps[j] = 0.0f;
  if (P.noise_mode == NOISE_EXTERNAL) {
    const float *nz = P.noise + ((size_t)c * P.noise_m + (size_t)(tg - P.noise_t0)) * D;
#pragma unroll
    for (int j = 0; j < D; ++j) eps[j] = __ldg(nz + j);
  } else if (P.noise_mode == NOISE_PHILOX) {
    noise_vec<D>(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, 0, eps);
  }
  stage_trajectory<D, 0, true>(P, c, pd, swin, sbase);

  // this part's run of waypoints (wpp is even: pairs never straddle two parts) and its
  // priors: the initial waypoint for both derivative passes at w = 0
  // (generate_trajectories.cu:285), else recomputed from the two waypoints before
  const uint32_t w0 = MP ? part * P.wpp : 0u;
  const uint32_t w_end = MP ? ((w0 + P.wpp) < P.n ? (w0 + P.wpp) : P.n) : P.n;
  float xprev[D], vprev[D];
  bool fine1 = P.fast_ok != 0, fine2 = fine1;
  if (!MP || w0 == 0) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      xprev[j] = vprev[j] = __ldg(P.start + j);
      fine1 = fine1 & fine_angle(xprev[j]);
    }
  } else if (w0 < P.n) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const float x1 = add_(sbase[j * NP + w0 - 1], eps[j]);
      const float x2 = add_(sbase[j * NP + w0 - 2], eps[j]); // w0 >= 2
      xprev[j] = x1;
      vprev[j] = div_fast(sub_(x1, x2), P.dt_div);
      fine1 = fine1 & fine_angle(x1);
      fine2 = fine2 & fine_angle(x2);
    }
  }

  float total = 0.0f, ltotal = 0.0f;
  int coll = 0;
  for (uint32_t w = w0; w < w_end; w += 2) {
    const bool has_hi = w + 1 < w_end;
    f2 x[D], vnew[D];
#pragma unroll
    for (int j = 0; j < D; ++j) { // generate_trajectories.cu:53-57
      const float2 b = *reinterpret_cast<const float2 *>(sbase + j * NP + w);
      x[j] = f2_pack(add_(b.x, eps[j]), add_(b.y, eps[j]));
    }
    f2 a2, l2;
    int cw0, cw1;
    bool f0, f1;
    waypoint_cost2<D, GM, LIM, FKM>(P, x, xprev, vprev, vnew, &a2, &l2, &cw0, &cw1, &f0, &f1);
    const bool valid0 = f0 & fine1 & fine2, valid1 = f1 & f0 & fine1;
    fine2 = f0 & (P.fast_ok != 0);
    fine1 = f1 & (P.fast_ok != 0);
    if (!(valid0 & valid1)) { // rare: redo the offending waypoint(s) with IEEE divides, full sin/cos
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        if (h == 0 ? valid0 : valid1) continue;
        WpIn<D> in;
#pragma unroll
        for (int j = 0; j < D; ++j) { // the hi waypoint's priors are the (possibly redone) lo results
          in.x[j] = h ? f2_hi(x[j]) : f2_lo(x[j]);
          in.xprev[j] = h ? f2_lo(x[j]) : xprev[j];
          in.vprev[j] = h ? f2_lo(vnew[j]) : vprev[j];
        }
        const WpOut<D> o = waypoint_cost_general<D, GM, true, LIM, FKM>(P, in);
#pragma unroll
        for (int j = 0; j < D; ++j)
          vnew[j] = h ? f2_pack(f2_lo(vnew[j]), o.vnew[j]) : f2_pack(o.vnew[j], f2_hi(vnew[j]));
        a2 = h ? f2_pack(f2_lo(a2), o.a2) : f2_pack(o.a2, f2_hi(a2));
        l2 = h ? f2_pack(f2_lo(l2), o.l2) : f2_pack(o.l2, f2_hi(l2));
        if (h) cw1 = o.coll; else cw0 = o.coll;
      }
    }
#pragma unroll
    for (int j = 0; j < D; ++j) {
      xprev[j] = f2_hi(x[j]);
      vprev[j] = f2_hi(vnew[j]);
    }
    coll += cw0;
    if (has_hi) coll += cw1;
    if (!MP) { // fixed order: left fold over the waypoints (oracle ORC_SUM_WAYPOINT)
      total = add_(total, f2_lo(a2));
      if (LIM) ltotal = add_(ltotal, f2_lo(l2));
      if (has_hi) {
        total = add_(total, f2_hi(a2));
        if (LIM) ltotal = add_(ltotal, f2_hi(l2));
      }
    } else { // slot w + 1 <= WS - 1 always exists
      float *sp = spart + ((size_t)swi * WS + w) * 32 + lane;
      sp[0] = f2_lo(a2);
      sp[32] = f2_hi(a2);
      if (LIM) {
        float *sl = slim + ((size_t)swi * WS + w) * 32 + lane;
        sl[0] = f2_lo(l2);
        sl[32] = f2_hi(l2);
      }
    }
  }
  if (MP) {
    scoll[warp * 32 + lane] = coll;
    __syncthreads();
    if (part != 0) return;
    coll = 0;
    for (uint32_t q = 0; q < T; ++q) coll += scoll[(swi * T + q) * 32 + lane];
    for (uint32_t ww = 0; ww < P.n; ++ww) { // the same left fold, waypoint 0 first
      total = add_(total, spart[((size_t)swi * WS + ww) * 32 + lane]);
      if (LIM) ltotal = add_(ltotal, slim[((size_t)swi * WS + ww) * 32 + lane]);
    }
  }

  float score = add_((float)coll, mul_(P.smooth_scale, total)); // :81,100,107-108
  if (LIM && P.limit_weight != 0.0f) score = add_(score, mul_(P.limit_weight, ltotal));
  if (active) {
    const size_t si = (size_t)c * P.m + t;
    if (P.scores) P.scores[si] = score;
    if (P.collisions) P.collisions[si] = coll;
  }
  if (P.keys_float) {
    unsigned long long key =
        active ? ((unsigned long long)key_hi_float(score) << 32) | tg : kKeyNone;
    key = warp_min_u64(key);
    if (lane == 0) atomicMin(P.keys_float + c, key);
  }
  if (P.keys) {
    unsigned long long key =
        active ? ((unsigned long long)key_hi(P.select_mode, score) << 32) | tg : kKeyNone;
    key = warp_min_u64(key);
    if (lane == 0) atomicMin(P.keys + c, key);
  }
}

// ------------------------------------------------------- update-vector kernel
// compute_update_vector (generate_trajectories.cu:117-133): u[j] = sum over the climber's
// samples of (best_score / ceilf(score_t)) * noise[t][j].  The reference sums with unordered
// float atomicAdd; here every term is rounded to a multiple of 2^-28 and summed with 64-bit
// integer atomics, so the result is independent of the order, the launch geometry and the
// number of GPUs (oracle: orc_update_sums).  One thread per sample, grid-stride.
template <int D>
__global__ void __launch_bounds__(256) update_vector_kernel(const __grid_constant__ UpdateParams P) {
  __shared__ long long s_part[8][8];
  const uint32_t c = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // the block's best_score: int(ceilf(.)) minimum over ALL ranks' samples (generate_trajectories.cu:109-111)
  const float best = (float)(int)(uint32_t)(__ldg(P.keys + c) >> 32);
  long long acc[D], poisoned = 0;
#pragma unroll
  for (int j = 0; j < D; ++j) acc[j] = 0;
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < P.m; t += gridDim.x * blockDim.x) {
    const float score = __ldg(P.scores + (size_t)c * P.m + t);
    const float wgt = dm::div_(best, ceilf(score)); // :128-129
    const uint32_t tg = (uint32_t)(P.sample_offset + t);
    float eps[D];
    if (P.noise_mode == NOISE_EXTERNAL) {
      const float *nz = P.noise + ((size_t)c * P.noise_m + (size_t)(tg - P.noise_t0)) * D;
#pragma unroll
      for (int j = 0; j < D; ++j) eps[j] = __ldg(nz + j);
    } else {
      dm::noise_vec<D>(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, 0, eps);
    }
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const float term = dm::mul_(wgt, eps[j]); // :130-131
      if (fabsf(term) < kUpdateTermMax)
        acc[j] += __float2ll_rn(dm::mul_(term, kUpdateScale));
      else
        poisoned += 1; // non-finite or huge: the update vector becomes NaN
    }
  }
  auto wsum = [&](long long v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(kFull, v, off);
    return v;
  };
#pragma unroll
  for (int j = 0; j < D; ++j) {
    const long long v = wsum(acc[j]);
    if (lane == 0) s_part[warp][j] = v;
  }
  {
    const long long v = wsum(poisoned);
    if (lane == 0) s_part[warp][7] = v;
  }
  __syncthreads();
  if (threadIdx.x < 8 && (threadIdx.x < D || threadIdx.x == 7)) {
    long long v = 0;
    for (uint32_t q = 0; q < (blockDim.x >> 5); ++q) v += s_part[q][threadIdx.x];
    if (v) atomicAdd((unsigned long long *)(P.usum + (size_t)c * 8 + threadIdx.x), (unsigned long long)v);
  }
}

// --------------------------------------------------------------- final kernel
// planning_state.cu:113-127: strict-< argmin over the climbers' scores (first
// wins), then the winner's trajectory and backward-difference velocities.
__global__ void final_kernel(const __grid_constant__ FinalParams P) {
  __shared__ unsigned long long s_best;
  if (threadIdx.x == 0) s_best = kKeyNone;
  __syncthreads();
  unsigned long long mine = kKeyNone;
  for (uint32_t c = threadIdx.x; c < P.k; c += blockDim.x) {
    float sc = P.scores[c];
    if (sc < INFINITY) { // `scores[i] < lowest_cost` with lowest_cost = +inf
      unsigned long long key = ((unsigned long long)__float_as_uint(sc) << 32) | c;
      mine = key < mine ? key : mine;
    }
  }
  if (mine != kKeyNone) atomicMin(&s_best, mine);
  __syncthreads();
  const unsigned long long best = s_best;
  const bool any = best != kKeyNone;
  const uint32_t c = any ? (uint32_t)best : 0u;
  if (threadIdx.x == 0) {
    P.out[0] = any ? __uint_as_float((uint32_t)(best >> 32)) : INFINITY;
    long long climber = (any && P.found_better[c]) ? (long long)c : -1ll;
    memcpy(P.out + 2, &climber, 8);
  }
  const float *traj = P.traj + (size_t)c * P.d * P.npad;
  float *o_traj = P.out + 4;
  float *o_vel = o_traj + (size_t)P.n * P.d;
  for (uint32_t i = threadIdx.x; i < P.n * P.d; i += blockDim.x) {
    const uint32_t w = i / P.d, j = i - w * P.d;
    float x = traj[(size_t)j * P.npad + w];
    float prior = w ? traj[(size_t)j * P.npad + w - 1] : P.start[j];
    o_traj[i] = x;
    o_vel[i] = dm::div_(dm::sub_(x, prior), P.dt);
  }
}

// ------------------------------------------------------------- scatter kernel
// OccupancyGrid::set(double,double,double,val) for every point of a cloud
// (occupancy_grid.cpp:47-58).  Four points per thread: three 16-byte coalesced loads bring 48
// contiguous bytes, the index math is that of the lookups (threshold compares + verified
// reciprocal divides when the grid has them, else grid_index), one byte store per point;
// equal values make concurrent stores to a voxel benign.
template <bool FAST>
__device__ __forceinline__ void scatter_point(const ScatterParams &P, float x, float y, float z) {
  if (FAST) {
    const GridDev &g = P.grid;
    const float ax = dm::add_(x, g.x_off), ay = dm::add_(y, g.y_off), az = dm::add_(z, 0.0f);
    const bool in = (__float_as_uint(ax) < g.tx_bits) & (__float_as_uint(ay) < g.ty_bits) &
                    (__float_as_uint(az) < g.tz_bits);
    const dm::FastDiv rd{g.res, g.res_rcp, 0.0f, 0.0f};
    const uint32_t ix = __float2uint_rz(dm::div_fast_unguarded(ax, rd));
    const uint32_t iy = __float2uint_rz(dm::div_fast_unguarded(ay, rd));
    const uint32_t iz = __float2uint_rz(dm::div_fast_unguarded(az, rd));
    if (in) P.cells[(iz * g.ny + iy) * g.nx + ix] = (uint8_t)P.val;
  } else {
    const long long idx = grid_index(P.grid, x, y, z);
    if (idx >= 0) P.cells[idx] = (uint8_t)P.val;
  }
}

template <bool FAST>
__global__ void __launch_bounds__(256) scatter_kernel(const __grid_constant__ ScatterParams P) {
  const uint64_t groups = P.num_points / 4; // xyz is 16-byte aligned (launch_scatter checks)
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const float4 *v = reinterpret_cast<const float4 *>(P.xyz);
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    const float4 a = __ldg(v + 3 * g), b = __ldg(v + 3 * g + 1), c = __ldg(v + 3 * g + 2);
    scatter_point<FAST>(P, a.x, a.y, a.z);
    scatter_point<FAST>(P, a.w, b.x, b.y);
    scatter_point<FAST>(P, b.z, b.w, c.x);
    scatter_point<FAST>(P, c.y, c.z, c.w);
  }
  const uint64_t t = groups * 4 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < P.num_points) // tail
    scatter_point<FAST>(P, __ldg(P.xyz + 3 * t), __ldg(P.xyz + 3 * t + 1), __ldg(P.xyz + 3 * t + 2));
}

// unaligned clouds: one point per thread
template <bool FAST>
__global__ void scatter_kernel_scalar(const __grid_constant__ ScatterParams P) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.num_points; i += stride)
    scatter_point<FAST>(P, __ldg(P.xyz + 3 * i), __ldg(P.xyz + 3 * i + 1), __ldg(P.xyz + 3 * i + 2));
}

// ------------------------------------------------- Euclidean distance transform
// Exact EDT of the occupancy grid (oracle: orc_distance_field; replaces the host KD-tree fill
// CPUDistanceGrid::update, cpu_distance_grid.cu:87-202).  Squared distances are integers in
// voxel units, so the result is bit-identical to the CPU restatement.
//
// Rows: one warp per row.  The row's occupancy becomes a bitmask, one 32-voxel word per lane
// (ballot); the nearest occupied voxel to the left / right of a voxel is a clz / ffs inside its
// word or a carry from the nearest non-empty word (warp scans).  Coalesced reads and writes.
__global__ void __launch_bounds__(128) edt_rows_kernel(const uint8_t *occ, uint32_t nx,
                                                       uint64_t rows, int32_t *out) {
  const uint64_t row = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const uint8_t *r = occ + row * nx;
  const uint32_t chunks = (nx + 31) / 32; // <= 32 (nx <= kEdtMaxDim)
  uint32_t mine = 0;                      // occupancy word of chunk `lane`
  for (uint32_t c = 0; c < chunks; ++c) {
    const uint32_t x = c * 32 + lane;
    const uint32_t m = __ballot_sync(kFull, x < nx && __ldg(r + x) != 0);
    if ((uint32_t)lane == c) mine = m;
  }
  // nearest occupied x in the chunks strictly before / after this lane's chunk
  int lastl = mine ? lane * 32 + 31 - __clz(mine) : -kEdtInf;
  int firstr = mine ? lane * 32 + __ffs(mine) - 1 : kEdtInf;
  int cl = lastl, cr = firstr;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int a = __shfl_up_sync(kFull, cl, off), b = __shfl_down_sync(kFull, cr, off);
    if (lane >= off) cl = max(cl, a);
    if (lane + off < 32) cr = min(cr, b);
  }
  int carry_l = __shfl_up_sync(kFull, cl, 1), carry_r = __shfl_down_sync(kFull, cr, 1);
  if (lane == 0) carry_l = -kEdtInf;
  if (lane == 31) carry_r = kEdtInf;
  for (uint32_t c = 0; c < chunks; ++c) {
    const uint32_t m = __shfl_sync(kFull, mine, c);
    const int kl = __shfl_sync(kFull, carry_l, c), kr = __shfl_sync(kFull, carry_r, c);
    const int x = (int)(c * 32) + lane;
    const uint32_t le = m & (0xFFFFFFFFu >> (31 - lane)), ge = m & (0xFFFFFFFFu << lane);
    const int xl = le ? (int)(c * 32) + 31 - __clz(le) : kl;
    const int xr = ge ? (int)(c * 32) + __ffs(ge) - 1 : kr;
    const int dl = xl < 0 ? kEdtInf : x - xl, dr = xr >= kEdtInf ? kEdtInf : xr - x;
    const int d = min(dl, dr);
    if ((uint32_t)x < nx) out[row * nx + x] = d >= kEdtInf ? kEdtInf : d * d;
  }
}

// Lines along y or z: one thread per line, the 32 lanes of the block on 32 consecutive x
// (coalesced), lower-envelope scan of Meijster et al. with the line's values and the two stacks
// in shared memory.  in/out element u of the line of lane l: base + u * stride + l.
template <bool TO_FLOAT>
__global__ void __launch_bounds__(32) edt_lines_kernel(const int32_t *in, void *out_v, uint32_t nx,
                                                       uint32_t n, uint64_t stride,
                                                       uint64_t line_stride, float res) {
  extern __shared__ __align__(16) unsigned char edt_smem[];
  int32_t *g = reinterpret_cast<int32_t *>(edt_smem);           // [n][32]
  uint16_t *s = reinterpret_cast<uint16_t *>(g + (size_t)n * 32); // [n][32]
  uint16_t *t = s + (size_t)n * 32;                              // [n][32]
  const int lane = threadIdx.x;
  const uint32_t x = blockIdx.x * 32 + lane;
  const bool live = x < nx;
  const uint64_t base = (uint64_t)blockIdx.y * line_stride + (live ? x : nx - 1);
  for (uint32_t u = 0; u < n; ++u) g[u * 32 + lane] = __ldg(in + base + (uint64_t)u * stride);
  auto f = [&](int xx, int i) -> long long {
    return (long long)(xx - i) * (xx - i) + g[i * 32 + lane];
  };
  int q = 0;
  s[lane] = 0;
  t[lane] = 0;
  for (int u = 1; u < (int)n; ++u) {
    while (q >= 0 && f(t[q * 32 + lane], s[q * 32 + lane]) > f(t[q * 32 + lane], u)) --q;
    if (q < 0) {
      q = 0;
      s[lane] = (uint16_t)u;
    } else {
      const long long i = s[q * 32 + lane];
      const long long num = (long long)u * u - i * i + g[u * 32 + lane] - g[i * 32 + lane];
      const long long den = 2 * (u - i);
      long long w = num / den;
      if ((num % den != 0) && (num < 0)) --w; // floor division
      w += 1;
      if (w < (long long)n) {
        ++q;
        s[q * 32 + lane] = (uint16_t)u;
        t[q * 32 + lane] = (uint16_t)w;
      }
    }
  }
  for (int u = (int)n - 1; u >= 0; --u) {
    const long long v = f(u, s[q * 32 + lane]);
    if (live) {
      if (TO_FLOAT)
        reinterpret_cast<float *>(out_v)[base + (uint64_t)u * stride] =
            v >= kEdtInf ? INFINITY : dm::mul_(sqrtf((float)v), res);
      else
        reinterpret_cast<int32_t *>(out_v)[base + (uint64_t)u * stride] = v >= kEdtInf ? kEdtInf : (int32_t)v;
    }
    if (u == (int)t[q * 32 + lane]) --q;
  }
}

// ------------------------------------------------------- depth -> voxel kernel
// One thread per pixel of every camera: deprojection exactly as convert_sensor_msg_to_pcl
// (mapping_node.cpp:110-135: x = (u - cx) * depth * (0.001 / fx), z = depth * 0.001f, depth 0
// skipped), rigid transform into the arm frame (:195; products summed left to right, then the
// translation), crop box (:262-263, inclusive like pcl::CropBox), then the voxel is set as in
// scatter_kernel.  The PCL VoxelGrid down-sampling of :190-191 has no counterpart: setting a
// voxel is idempotent, so duplicates need no thinning.  Reads are 2-byte coalesced.
__global__ void depth_scatter_kernel(const __grid_constant__ DepthParams P) {
  const uint64_t per_cam = (uint64_t)P.w * P.h, total = per_cam * P.cams;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long kept = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const uint32_t cam = (uint32_t)(i / per_cam);
    const uint32_t pix = (uint32_t)(i - (uint64_t)cam * per_cam);
    const uint32_t v = pix / P.w, u = pix - v * P.w;
    const uint16_t d = __ldg(P.depth + i);
    if (d == 0) continue;
    const DepthCam &C = P.cam[cam];
    const float df = (float)d;
    const float x = dm::mul_(dm::mul_(dm::sub_((float)u, C.cx), df), C.kx);
    const float y = dm::mul_(dm::mul_(dm::sub_((float)v, C.cy), df), C.ky);
    const float z = dm::mul_(df, 0.001f);
    float q[3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
      q[a] = dm::add_(dm::add_(dm::add_(dm::mul_(C.r[3 * a], x), dm::mul_(C.r[3 * a + 1], y)),
                               dm::mul_(C.r[3 * a + 2], z)),
                      C.t[a]);
    const bool in = (q[0] >= P.crop_lo[0]) & (q[0] <= P.crop_hi[0]) & (q[1] >= P.crop_lo[1]) &
                    (q[1] <= P.crop_hi[1]) & (q[2] >= P.crop_lo[2]) & (q[2] <= P.crop_hi[2]);
    if (!in) continue;
    kept += 1;
    const long long idx = grid_index(P.grid, q[0], q[1], q[2]);
    if (idx >= 0) P.cells[idx] = (uint8_t)P.val;
  }
  if (P.kept) {
    kept = __reduce_add_sync(kFull, (unsigned)kept);
    if ((threadIdx.x & 31) == 0 && kept) atomicAdd(P.kept, kept);
  }
}

// ------------------------------------------------------------ gather probe
// Roofline denominator for the collision lookups: independent random loads of one
// cell each (one 32-byte sector per load) from an array of a chosen size, 8 in flight per
// thread.  An 8 MiB array measures the L2 gather rate, one larger than L2 the HBM rate.
template <typename T>
__global__ void __launch_bounds__(256) gather_probe_kernel(const T *cells, uint64_t num_cells,
                                                           uint32_t per_thread, uint32_t *sink) {
  uint32_t h = (blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B9u + 0x7F4A7C15u;
  uint32_t acc = 0;
  for (uint32_t i = 0; i < per_thread; i += 8) {
    uint64_t idx[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      h ^= h << 13; h ^= h >> 17; h ^= h << 5; // xorshift32
      idx[u] = ((uint64_t)h * num_cells) >> 32;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += (uint32_t)__ldg(cells + idx[u]);
  }
  if (acc == 0xFFFFFFFFu) *sink = acc; // keeps the loads alive
}

__global__ void lookup_kernel(GridDev g, const float *xyz, uint64_t count, float coll_dist,
                              long long *cell_index, uint8_t *collides) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  long long idx = grid_index(g, x, y, z);
  cell_index[i] = idx;
  collides[i] = (g.kind != 0 && g.cells) ? (uint8_t)grid_collides(g, x, y, z, coll_dist) : 0;
}

// noise[k][m][n_noise][d] exactly as the fused kernels draw it
__global__ void noise_fill_kernel(const __grid_constant__ CostParams P, int d, float *out) {
  const uint64_t total = (uint64_t)P.k * P.m * P.n_noise * d;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  uint32_t j = (uint32_t)(i % d);
  uint64_t r = i / d;
  uint32_t wn = (uint32_t)(r % P.n_noise);
  r /= P.n_noise;
  uint32_t t = (uint32_t)(r % P.m);
  uint32_t c = (uint32_t)(r / P.m);
  out[i] = dm::noise_at(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c,
                        (uint32_t)(P.sample_offset + t), wn, j);
}

// [k or 1][n][d] array-of-structs -> [k][d][npad] structure-of-arrays
__global__ void aos_to_soa_kernel(const float *aos, float *soa, uint32_t k, uint32_t n,
                                  uint32_t npad, uint32_t d, int broadcast_first) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t total = (uint64_t)k * d * npad;
  if (i >= total) return;
  uint32_t w = (uint32_t)(i % npad);
  uint32_t j = (uint32_t)((i / npad) % d);
  uint32_t c = (uint32_t)(i / ((uint64_t)npad * d));
  uint32_t cs = broadcast_first ? 0 : c;
  soa[i] = w < n ? aos[((size_t)cs * n + w) * d + j] : 0.0f;
}

__global__ void debug_sincos_kernel(const float *x, uint64_t count, float *s, float *c) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) dm::sincos(x[i], s + i, c + i);
}

template <int D>
__global__ void debug_fk_model_kernel(const FkModelDev *model, const float *wp, uint64_t count,
                                      uint32_t num_spheres, float *out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float th[D];
#pragma unroll
  for (int j = 0; j < D; ++j) th[j] = wp[i * D + j];
  fk_model_walk<D, false>(*model, th, [&](uint32_t q, P3 sp) {
    float *o = out + (i * num_spheres + model->order[q]) * 3;
    o[0] = sp.x;
    o[1] = sp.y;
    o[2] = sp.z;
  });
}

template <int D>
__global__ void debug_fk_kernel(const float *wp, uint64_t count, float rpy_s, float rpy_c,
                                float *out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float th[D];
#pragma unroll
  for (int j = 0; j < D; ++j) th[j] = wp[i * D + j];
  P3 pts[5];
  fk_points<D, false>(th, rpy_s, rpy_c, pts);
  for (int l = 0; l < 5; ++l) {
    out[(i * 5 + l) * 3 + 0] = pts[l].x;
    out[(i * 5 + l) * 3 + 1] = pts[l].y;
    out[(i * 5 + l) * 3 + 2] = pts[l].z;
  }
}

// Exhaustive check of the reciprocal-multiply division against the IEEE divide for every
// float whose bit pattern lies in [lo_bits, hi_bits].  mode 0: quotient bits must match;
// mode 1: the truncated quotient (voxel index) must match.
__global__ void verify_div_kernel(dm::FastDiv d, uint32_t lo_bits, uint32_t hi_bits, int mode,
                                  unsigned long long *bad) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long mine = 0;
  for (uint64_t u = (uint64_t)lo_bits + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       u <= hi_bits; u += stride) {
    const float a = __uint_as_float((uint32_t)u);
    const float f = dm::div_fast_unguarded(a, d), e = __fdiv_rn(a, d.b);
    if (mode == 0)
      mine += __float_as_uint(f) != __float_as_uint(e);
    else
      mine += __float2uint_rz(f) != __float2uint_rz(e);
  }
  mine = __reduce_add_sync(kFull, (unsigned)mine);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(bad, mine);
}

inline unsigned blocks_for(uint64_t total, unsigned threads) {
  return (unsigned)((total + threads - 1) / threads);
}

} // namespace

void host_sincos(float x, float *s, float *c) { dm::sincos(x, s, c); }

template <int D, int GM, bool NPW, bool LIM, bool FKM = false>
static cudaError_t launch_cost_t(const CostParams &p, dim3 grid, unsigned threads, size_t smem,
                                 unsigned cluster, cudaStream_t st) {
  // the specialised CDG kernels exist only with the verified reciprocal index divide; a grid
  // whose verification failed (or DOAPP_EXACT_DIV=1) runs the generic kernel (IEEE divides)
  constexpr bool FIDX = GM == GM_U8_CDG || GM == GM_F32_CDG;
  auto kern = cost_kernel<D, GM, NPW, FIDX, LIM, FKM>;
  if (smem > 48 * 1024) {
    cudaError_t e =
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = cluster > 1 ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int D, int GM, bool LIM, bool FKM, bool MP>
static cudaError_t launch_cost_pk_m(const CostParams &p, cudaStream_t st) {
  const uint32_t T = 1u << p.parts_log2, SW = (kPkThreads / 32) >> p.parts_log2;
  size_t smem = ((size_t)D * (p.npad + 2) + 8) * sizeof(float);
  if (MP) smem += ((size_t)SW * T * p.wpp * 32 * (LIM ? 2 : 1) + 4 * 32) * sizeof(float);
  auto kern = cost_kernel_pk<D, GM, LIM, FKM, MP>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const uint32_t spb = SW * 32; // samples per CTA
  const dim3 grid((p.m + spb - 1) / spb, p.k);
  kern<<<grid, kPkThreads, smem, st>>>(p);
  return cudaGetLastError();
}

template <int D, int GM, bool LIM, bool FKM = false>
static cudaError_t launch_cost_pk_t(const CostParams &p, cudaStream_t st) {
  return p.parts_log2 ? launch_cost_pk_m<D, GM, LIM, FKM, true>(p, st)
                      : launch_cost_pk_m<D, GM, LIM, FKM, false>(p, st);
}

template <int D>
static cudaError_t launch_cost_pk_d(const CostParams &p, int gm, cudaStream_t st) {
  const bool lim = p.use_limits != 0;
#define DOAPP_PK(GMV)                                                                 \
  return lim ? launch_cost_pk_t<D, GMV, true>(p, st) : launch_cost_pk_t<D, GMV, false>(p, st)
#define DOAPP_PK_FK(GMV)                                                              \
  return lim ? launch_cost_pk_t<D, GMV, true, true>(p, st) : launch_cost_pk_t<D, GMV, false, true>(p, st)
  if (p.fk_on) { // URDF chain + spheres
    if (gm == GM_U8_CDG) DOAPP_PK_FK(GM_U8_CDG);
    if (gm == GM_F32_CDG) DOAPP_PK_FK(GM_F32_CDG);
  }
  switch (gm) {
  case GM_U8_CDG: DOAPP_PK(GM_U8_CDG);
  case GM_F32_CDG: DOAPP_PK(GM_F32_CDG);
  default: DOAPP_PK(GM_NONE);
  }
#undef DOAPP_PK
#undef DOAPP_PK_FK
}

static int grid_mode_of(const CostParams &p) {
  int gm = GM_NONE;
  if (p.collision_mode != 0 && p.grid.kind != 0 && p.grid.cells) {
    gm = GM_GENERIC;
    if (p.grid.frame == 0 && p.grid.fast_idx) gm = p.grid.kind == 1 ? GM_U8_CDG : GM_F32_CDG;
  }
  return gm;
}

// ---- persistent planner: kernel selection, geometry, cooperative launch
template <int D>
static const void *persistent_kernel_of(int gm, bool lim, bool fkm) {
#define DOAPP_PP(GMV)                                                    \
  return lim ? (const void *)plan_persistent_kernel<D, GMV, true, false> \
             : (const void *)plan_persistent_kernel<D, GMV, false, false>
#define DOAPP_PP_FK(GMV)                                                \
  return lim ? (const void *)plan_persistent_kernel<D, GMV, true, true> \
             : (const void *)plan_persistent_kernel<D, GMV, false, true>
  if (fkm) { // URDF chain + spheres: own instances (register budget of the reference chain kept)
    switch (gm) {
    case GM_U8_CDG: DOAPP_PP_FK(GM_U8_CDG);
    case GM_F32_CDG: DOAPP_PP_FK(GM_F32_CDG);
    case GM_GENERIC: DOAPP_PP_FK(GM_GENERIC);
    default: break;
    }
  }
  switch (gm) {
  case GM_U8_CDG: DOAPP_PP(GM_U8_CDG);
  case GM_F32_CDG: DOAPP_PP(GM_F32_CDG);
  case GM_GENERIC: DOAPP_PP(GM_GENERIC);
  default: DOAPP_PP(GM_NONE);
  }
#undef DOAPP_PP
#undef DOAPP_PP_FK
}

static size_t persistent_smem(const CostParams &p, int d, uint32_t T, uint32_t wpp) {
  const uint32_t SB = 256 / T, WS = T * wpp;
  size_t f = (size_t)d * p.npad + (size_t)SB * WS * (p.use_limits ? 2 : 1) + (size_t)SB * 8 + 8 +
             ((SB + 1) & ~1u) + 8 + (size_t)SB * 8;
  return f * sizeof(float);
}

cudaError_t persistent_plan_config(const CostParams &p, int d, int sm_count, PersistParams *q,
                                   bool *ok) {
  *ok = false;
  if (p.k == 0 || p.m == 0 || p.n_noise != 1) return cudaSuccess;
  const void *kern = d == 5 ? persistent_kernel_of<5>(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0)
                            : persistent_kernel_of<4>(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0);
  // fewest waypoints per thread (shortest dependent chain) whose grid is co-resident
  for (uint32_t wpp = 1; wpp <= 4; wpp *= 2) {
    const uint32_t need = (p.n + wpp - 1) / wpp;
    uint32_t tl = 0;
    while ((1u << tl) < need) ++tl;
    if (tl > 8) continue;
    const uint32_t T = 1u << tl, SB = 256 / T;
    const size_t smem = persistent_smem(p, d, T, wpp);
    if (smem > 200 * 1024) continue;
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kPersistThreads, smem);
    if (e != cudaSuccess) return e;
    const uint64_t ctas = (uint64_t)p.k * ((p.m + SB - 1) / SB);
    if (ctas > (uint64_t)per_sm * sm_count) continue;
    q->cost = p;
    q->cost.res_div = dm::FastDiv{p.grid.res, p.grid.res_rcp, 0.0f, 0.0f};
    q->cost.wpp = wpp;
    q->t_log2 = tl;
    *ok = true;
    return cudaSuccess;
  }
  return cudaSuccess;
}

cudaError_t launch_persistent_plan(const PersistParams &q, int d, cudaStream_t st) {
  const CostParams &p = q.cost;
  const void *kern = d == 5 ? persistent_kernel_of<5>(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0)
                            : persistent_kernel_of<4>(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0);
  const uint32_t T = 1u << q.t_log2, SB = 256 / T;
  const size_t smem = persistent_smem(p, d, T, p.wpp);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  void *args[] = {(void *)&q};
  return cudaLaunchCooperativeKernel(kern, dim3((p.m + SB - 1) / SB, p.k), dim3(kPersistThreads), args, smem, st);
}

template <int D>
static cudaError_t launch_cost_d(const CostParams &p, dim3 grid, unsigned threads, size_t smem,
                                 unsigned cluster, cudaStream_t st) {
  const int gm = grid_mode_of(p);
  const bool npw = p.n_noise != 1 && p.noise_mode != NOISE_ZERO;
  const bool lim = p.use_limits != 0;
#define DOAPP_GO(GMV)                                                                       \
  return npw ? (lim ? launch_cost_t<D, GMV, true, true>(p, grid, threads, smem, cluster, st)  \
                    : launch_cost_t<D, GMV, true, false>(p, grid, threads, smem, cluster, st)) \
             : (lim ? launch_cost_t<D, GMV, false, true>(p, grid, threads, smem, cluster, st) \
                    : launch_cost_t<D, GMV, false, false>(p, grid, threads, smem, cluster, st))
#define DOAPP_GO_FK(GMV)                                                                          \
  return npw ? (lim ? launch_cost_t<D, GMV, true, true, true>(p, grid, threads, smem, cluster, st)  \
                    : launch_cost_t<D, GMV, true, false, true>(p, grid, threads, smem, cluster, st)) \
             : (lim ? launch_cost_t<D, GMV, false, true, true>(p, grid, threads, smem, cluster, st) \
                    : launch_cost_t<D, GMV, false, false, true>(p, grid, threads, smem, cluster, st))
  if (p.fk_on) { // URDF chain + spheres (doapp_set_fk_model): its own kernel instances
    switch (gm) {
    case GM_U8_CDG: DOAPP_GO_FK(GM_U8_CDG);
    case GM_F32_CDG: DOAPP_GO_FK(GM_F32_CDG);
    case GM_GENERIC: DOAPP_GO_FK(GM_GENERIC);
    default: break; // no grid: no collision term, either chain
    }
  }
  switch (gm) {
  case GM_U8_CDG: DOAPP_GO(GM_U8_CDG);
  case GM_F32_CDG: DOAPP_GO(GM_F32_CDG);
  case GM_GENERIC: DOAPP_GO(GM_GENERIC);
  default: DOAPP_GO(GM_NONE);
  }
#undef DOAPP_GO
#undef DOAPP_GO_FK
}

cudaError_t launch_cost(const CostParams &p_in, int d, int sm_count, cudaStream_t st) {
  CostParams p = p_in;
  if (p.k == 0 || p.m == 0) return cudaSuccess;
  p.res_div = dm::FastDiv{p.grid.res, p.grid.res_rcp, 0.0f, 0.0f};
  // Threads per sample: 1 when the samples alone fill the chip; otherwise each sample's
  // waypoints are split over T = 2..32 threads (4 warps per CTA, T/4 CTAs per cluster) so
  // that a small problem (config 2: 1024 samples x 64 waypoints) still spreads over every
  // SM and the dependent chain per thread is a couple of waypoints.
  const int dp = d == 4 ? 4 : 8;
  const uint64_t sample_warps = (uint64_t)p.k * ((p.m + 31) / 32);
  const uint64_t target_warps = (uint64_t)sm_count * 12;
  auto smem_for = [&](uint32_t T) -> size_t {
    size_t f = (size_t)p.n * dp + 8;
    if (T >= 4) f += (size_t)d * 32;
    if (T > 1) f += (size_t)p.n * 32 * (p.use_limits ? 2 : 1) + (size_t)T * 32;
    return f * sizeof(float);
  };
  uint32_t lg = 0;
  while (lg < 5 && (2u << lg) <= p.n && sample_warps * (1u << lg) < target_warps &&
         smem_for(2u << lg) <= 200 * 1024)
    ++lg;
  {
    // tuning override: DOAPP_PARTS_LOG2 caps the threads per sample (profiling only)
    static const char *env = getenv("DOAPP_PARTS_LOG2");
    if (env && env[0] >= '0' && env[0] <= '5' && (uint32_t)(env[0] - '0') < lg) lg = env[0] - '0';
  }
  {
    // One thread per sample with two waypoints per step in packed fp32x2 arithmetic whenever
    // the samples alone fill the chip (DOAPP_PK=0/1 forces the choice where it is legal).
    const bool npw = p.n_noise != 1 && p.noise_mode != NOISE_ZERO;
    const int gm = grid_mode_of(p);
    // Up to 4 parts per sample stay inside one CTA of the packed kernel (its per-waypoint sums
    // live in shared memory: n <= kPkMaxPartN); more parts need the scalar kernel's clusters.
    uint32_t plg = lg <= 2 ? lg : 2;
    if (p.n > kPkMaxPartN || p.n < 4) plg = 0;
    bool pk = lg == plg;
    if (p.pk_mode == 1) pk = true;
    if (p.pk_mode == 0) pk = false;
    if (pk && !npw && gm != GM_GENERIC && p.fast_ok) {
      p.parts_log2 = plg;
      const uint32_t per = (p.n + (1u << plg) - 1) >> plg; // waypoints per part, rounded up to even
      p.wpp = (per + 1) & ~1u;
      return d == 5 ? launch_cost_pk_d<5>(p, gm, st) : launch_cost_pk_d<4>(p, gm, st);
    }
  }
  p.parts_log2 = lg;
  const uint32_t T = 1u << lg;
  p.pw_log2 = lg < 2 ? lg : 2;
  const uint32_t PW = 1u << p.pw_log2, cs = T / PW;
  p.wpp = (p.n + T - 1) / T;
  const uint32_t sw = T > 1 ? 1 : 4; // sample-warps per CTA
  const unsigned threads = 32 * PW * sw;
  const uint32_t spb = 32 * sw;
  const dim3 grid(((p.m + spb - 1) / spb) * cs, p.k);
  const size_t smem = smem_for(T);
  return d == 5 ? launch_cost_d<5>(p, grid, threads, smem, cs, st)
                : launch_cost_d<4>(p, grid, threads, smem, cs, st);
}

cudaError_t launch_verify_div(dm::FastDiv d, uint32_t lo_bits, uint32_t hi_bits, int mode,
                              unsigned long long *bad, int sm_count, cudaStream_t st) {
  verify_div_kernel<<<sm_count * 8, 256, 0, st>>>(d, lo_bits, hi_bits, mode, bad);
  return cudaGetLastError();
}

cudaError_t launch_update_vector(const UpdateParams &p, int sm_count, cudaStream_t st) {
  if (p.k == 0 || p.m == 0) return cudaSuccess;
  unsigned bx = blocks_for(p.m, 256);
  const unsigned cap = (unsigned)((sm_count * 8 + p.k - 1) / p.k);
  if (bx > cap) bx = cap > 0 ? cap : 1;
  const dim3 grid(bx, p.k);
  if (p.d == 5)
    update_vector_kernel<5><<<grid, 256, 0, st>>>(p);
  else
    update_vector_kernel<4><<<grid, 256, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_final(const FinalParams &p, cudaStream_t st) {
  final_kernel<<<1, 256, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_distance_field(const uint8_t *occ, uint32_t nx, uint32_t ny, uint32_t nz,
                                  float res, int32_t *tmp_a, int32_t *tmp_b, float *out,
                                  cudaStream_t st) {
  const uint64_t rows = (uint64_t)ny * nz;
  edt_rows_kernel<<<(unsigned)((rows + 3) / 4), 128, 0, st>>>(occ, nx, rows, tmp_a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  auto smem_of = [](uint32_t n) { return (size_t)n * 32 * (4 + 2 + 2); };
  const size_t sy = smem_of(ny), sz = smem_of(nz);
  if (sy > 48 * 1024 &&
      (e = cudaFuncSetAttribute(edt_lines_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sy)) != cudaSuccess)
    return e;
  if (sz > 48 * 1024 &&
      (e = cudaFuncSetAttribute(edt_lines_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sz)) != cudaSuccess)
    return e;
  // along y: lines of plane z, element stride nx
  edt_lines_kernel<false><<<dim3((nx + 31) / 32, nz), 32, sy, st>>>(tmp_a, tmp_b, nx, ny, nx,
                                                                  (uint64_t)nx * ny, res);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  // along z: lines of row y, element stride nx * ny
  edt_lines_kernel<true><<<dim3((nx + 31) / 32, ny), 32, sz, st>>>(tmp_b, out, nx, nz,
                                                                 (uint64_t)nx * ny, nx, res);
  return cudaGetLastError();
}

cudaError_t launch_depth_scatter(const DepthParams &p, int sm_count, cudaStream_t st) {
  const uint64_t total = (uint64_t)p.w * p.h * p.cams;
  if (total == 0) return cudaSuccess;
  unsigned blocks = blocks_for(total, 256);
  const unsigned cap = (unsigned)sm_count * 8;
  if (blocks > cap) blocks = cap;
  depth_scatter_kernel<<<blocks, 256, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_scatter(const ScatterParams &p, int sm_count, cudaStream_t st) {
  if (p.num_points == 0) return cudaSuccess;
  const bool fast = p.grid.frame == 0 && p.grid.fast_idx;
  const bool aligned = ((uintptr_t)p.xyz & 15u) == 0;
  const unsigned cap = (unsigned)sm_count * 8;
  if (aligned) {
    unsigned blocks = blocks_for((p.num_points + 3) / 4, 256);
    if (blocks > cap) blocks = cap;
    if (fast)
      scatter_kernel<true><<<blocks, 256, 0, st>>>(p);
    else
      scatter_kernel<false><<<blocks, 256, 0, st>>>(p);
  } else {
    unsigned blocks = blocks_for(p.num_points, 256);
    if (blocks > cap) blocks = cap;
    if (fast)
      scatter_kernel_scalar<true><<<blocks, 256, 0, st>>>(p);
    else
      scatter_kernel_scalar<false><<<blocks, 256, 0, st>>>(p);
  }
  return cudaGetLastError();
}

cudaError_t launch_gather_probe(const void *cells, uint64_t num_cells, uint32_t cell_bytes,
                                uint32_t blocks, uint32_t per_thread, uint32_t *sink,
                                cudaStream_t st) {
  if (cell_bytes == 4)
    gather_probe_kernel<float><<<blocks, 256, 0, st>>>((const float *)cells, num_cells, per_thread, sink);
  else
    gather_probe_kernel<uint8_t><<<blocks, 256, 0, st>>>((const uint8_t *)cells, num_cells, per_thread, sink);
  return cudaGetLastError();
}

cudaError_t launch_lookup(const GridDev &g, const float *xyz, uint64_t count, float coll_dist,
                          long long *cell_index, uint8_t *collides, cudaStream_t st) {
  if (count == 0) return cudaSuccess;
  lookup_kernel<<<blocks_for(count, 256), 256, 0, st>>>(g, xyz, count, coll_dist, cell_index,
                                                        collides);
  return cudaGetLastError();
}

cudaError_t launch_noise_fill(const CostParams &p, int d, float *out, cudaStream_t st) {
  uint64_t total = (uint64_t)p.k * p.m * p.n_noise * d;
  if (total == 0) return cudaSuccess;
  noise_fill_kernel<<<blocks_for(total, 256), 256, 0, st>>>(p, d, out);
  return cudaGetLastError();
}

cudaError_t launch_aos_to_soa(const float *aos, float *soa, uint32_t k, uint32_t n, uint32_t npad,
                              uint32_t d, int broadcast_first, cudaStream_t st) {
  uint64_t total = (uint64_t)k * d * npad;
  if (total == 0) return cudaSuccess;
  aos_to_soa_kernel<<<blocks_for(total, 256), 256, 0, st>>>(aos, soa, k, n, npad, d,
                                                            broadcast_first);
  return cudaGetLastError();
}

cudaError_t launch_debug_sincos(const float *x, uint64_t count, float *s, float *c,
                                cudaStream_t st) {
  if (count == 0) return cudaSuccess;
  debug_sincos_kernel<<<blocks_for(count, 256), 256, 0, st>>>(x, count, s, c);
  return cudaGetLastError();
}

cudaError_t launch_debug_fk_model(const FkModelDev *model, const float *wp, uint64_t count, int d,
                                  uint32_t num_spheres, float *out, cudaStream_t st) {
  if (count == 0 || num_spheres == 0) return cudaSuccess;
  if (d == 5)
    debug_fk_model_kernel<5><<<blocks_for(count, 128), 128, 0, st>>>(model, wp, count, num_spheres, out);
  else
    debug_fk_model_kernel<4><<<blocks_for(count, 128), 128, 0, st>>>(model, wp, count, num_spheres, out);
  return cudaGetLastError();
}

cudaError_t launch_debug_fk(const float *wp, uint64_t count, int d, float rpy_s, float rpy_c,
                            float *out, cudaStream_t st) {
  if (count == 0) return cudaSuccess;
  if (d == 5)
    debug_fk_kernel<5><<<blocks_for(count, 128), 128, 0, st>>>(wp, count, rpy_s, rpy_c, out);
  else
    debug_fk_kernel<4><<<blocks_for(count, 128), 128, 0, st>>>(wp, count, rpy_s, rpy_c, out);
  return cudaGetLastError();
}

} // namespace doapp
