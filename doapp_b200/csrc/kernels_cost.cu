// kernels_cost.cu (part of the kernel set described below; compiled once per joint count,
// -DDOAPP_D=4 and -DDOAPP_D=5) — hand-written sm_100a kernels of the perturbator hot path.
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
#include "cost_device.cuh"

#ifndef DOAPP_D
#error "compile with -DDOAPP_D=4 or -DDOAPP_D=5"
#endif

namespace doapp {

namespace {

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

} // namespace

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

cudaError_t DOAPP_DIM_FN(launch_cost_dim)(const CostParams &p, dim3 grid, unsigned threads,
                                          size_t smem, unsigned cluster, cudaStream_t st) {
  return launch_cost_d<DOAPP_D>(p, grid, threads, smem, cluster, st);
}

} // namespace doapp
