// cost_device.cuh — device functions shared by the cost / planner kernels (one copy per
// translation unit): grid lookup, the reference's FK chain and the URDF model, one waypoint's
// cost in scalar and packed (fp32x2) form, the pending keep-best prologue and the trajectory
// staging.  Included by kernels_cost.cu, kernels_pk.cu, kernels_persist.cu and kernels_aux.cu.
#pragma once

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
    unsigned long long pk;
    if (P.xr_nranks > 1 && P.pend_mode == PEND_KEEP_BEST) {
      // the ranks' keys of the previous iteration arrive in this rank's mailbox (local memory,
      // written by the peers over NVLink): lanes of warp 0 poll one rank each
      __shared__ unsigned long long s_pk;
      if (threadIdx.x < 32) {
        unsigned long long v = kKeyNone;
        if (threadIdx.x < P.xr_nranks) {
          const unsigned long long *slot = P.xr_recv + (size_t)c * 8 + threadIdx.x;
          unsigned long long t0 = 0;
          for (uint32_t spins = 1;; ++spins) {
            asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
            if (v != kKeyNone) break;
            if ((spins & 0x3FFu) == 0) {
              unsigned long long now;
              asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
              if (t0 == 0) t0 = now;
              if (now - t0 > P.xr_timeout_ns || *(volatile unsigned int *)P.xr_error != 0) {
                atomicExch(P.xr_error, 1u);
                break;
              }
            }
          }
        }
        v = warp_min_u64(v);
        if (threadIdx.x == 0) s_pk = v;
      }
      __syncthreads();
      pk = s_pk;
    } else {
      pk = __ldg(P.pend_keys + c);
    }
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

// Raw cell values of two points (threshold form, see GridDev::tx_bits): the occupancy byte or the
// bits of the distance; a point outside the grid reads as "no hit" (0 / +inf).  Separating the
// load from the predicate lets a kernel count the hits LATER (cost_kernel_pk counts a pair's
// lookups one iteration after it issued them, so an L2 round trip hides behind arithmetic).
template <int GM> __device__ __forceinline__ uint32_t raw_no_hit() {
  return GM == GM_U8_CDG ? 0u : 0x7f800000u;
}
template <int GM> __device__ __forceinline__ int hit_of(uint32_t raw, float thr) {
  if (GM == GM_U8_CDG) return raw != 0u;
  return __uint_as_float(raw) <= thr; // generate_trajectories.cu:412
}
template <int GM>
__device__ __forceinline__ void lookup2(const CostParams &P, P3v p, uint32_t *raw_lo, uint32_t *raw_hi) {
  using namespace dm;
  const GridDev &g = P.grid;
  const f2 ax = f2_add(p.x, f2_splat(g.x_off)), ay = f2_add(p.y, f2_splat(g.y_off));
  const f2 az = f2_add(p.z, f2_splat(0.0f));
  const f2 fx = f2_div_fast_unguarded(ax, P.res_div), fy = f2_div_fast_unguarded(ay, P.res_div),
           fz = f2_div_fast_unguarded(az, P.res_div);
  auto half = [&](float axh, float ayh, float azh, float fxh, float fyh, float fzh) -> uint32_t {
    const bool in = (__float_as_uint(axh) < g.tx_bits) & (__float_as_uint(ayh) < g.ty_bits) &
                    (__float_as_uint(azh) < g.tz_bits);
    const uint32_t ix = __float2uint_rz(fxh), iy = __float2uint_rz(fyh), iz = __float2uint_rz(fzh);
    const uint32_t lin = (iz * g.ny + iy) * g.nx + ix;
    uint32_t raw = raw_no_hit<GM>();
    if (GM == GM_U8_CDG) {
      if (in) raw = __ldg((const uint8_t *)g.cells + lin);
    } else {
      if (in) raw = __float_as_uint(__ldg((const float *)g.cells + lin));
    }
    return raw;
  };
  *raw_lo = half(f2_lo(ax), f2_lo(ay), f2_lo(az), f2_lo(fx), f2_lo(fy), f2_lo(fz));
  *raw_hi = half(f2_hi(ax), f2_hi(ay), f2_hi(az), f2_hi(fx), f2_hi(fy), f2_hi(fz));
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

// collision counts of two waypoints (lo, hi): the reference chain's 5 link points or the FK
// model's spheres
template <int D, int GM, bool FKM>
__device__ __forceinline__ void waypoint_collisions2(const CostParams &P, const dm::f2 (&x)[D],
                                                     int *coll_lo, int *coll_hi) {
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
  *coll_lo = c_lo;
  *coll_hi = c_hi;
}

// the reference chain's five link points of two waypoints: raw cell values, to be counted later
template <int D, int GM>
__device__ __forceinline__ void waypoint_lookups2(const CostParams &P, const dm::f2 (&x)[D],
                                                  uint32_t (&raw_lo)[5], uint32_t (&raw_hi)[5]) {
  P3v pts[5];
  fk_points2<D>(x, P.rpy_s, P.rpy_c, pts);
#pragma unroll
  for (int l = 0; l < 5; ++l) lookup2<GM>(P, pts[l], &raw_lo[l], &raw_hi[l]);
}

// waypoint_cost<..., FAST = true> on two CONSECUTIVE WAYPOINTS (lo = w, hi = w + 1) of one
// sample.  xprev / vprev are waypoint w - 1's angles and velocities; the hi half's priors are
// the lo half's own results.  fine_lo / fine_hi as waypoint_cost's return value.
template <int D, int GM, bool LIM, bool FKM, bool COLL = true>
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
  if (COLL) {
    waypoint_collisions2<D, GM, FKM>(P, x, coll_lo, coll_hi);
  } else { // the caller looks the link points up itself (waypoint_lookups2)
    *coll_lo = 0;
    *coll_hi = 0;
  }
  *a2_out = a2;
  *l2_out = l2;
  *fine_lo = f_lo;
  *fine_hi = f_hi;
}

inline unsigned blocks_for(uint64_t total, unsigned threads) {
  return (unsigned)((total + threads - 1) / threads);
}

inline int grid_mode_of(const CostParams &p) {
  int gm = GM_NONE;
  if (p.collision_mode != 0 && p.grid.kind != 0 && p.grid.cells) {
    gm = GM_GENERIC;
    if (p.grid.frame == 0 && p.grid.fast_idx) gm = p.grid.kind == 1 ? GM_U8_CDG : GM_F32_CDG;
  }
  return gm;
}

} // namespace

} // namespace doapp
