// kernels_pk.cu — cost_kernel_pk: the cost pass with one thread per sample (or 2 / 4 parts),
// two waypoints per step in packed fp32x2 arithmetic (see kernels_cost.cu for the stage map).
// Compiled once per joint count (-DDOAPP_D=4 / 5).
#include "cost_device.cuh"

#ifndef DOAPP_D
#error "compile with -DDOAPP_D=4 or -DDOAPP_D=5"
#endif

namespace doapp {

namespace {

constexpr int kPkThreads = 128;
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
  bool bad = false;
  // Reference chain on a grid: the ten cell loads of a pair are issued in its iteration and
  // COUNTED in the next one, so that an L1 miss (the 512^3 grid of BASELINE config 5 hits L1
  // only 55 % of the time) waits behind a whole pair's arithmetic instead of ~60 instructions.
  // Multi-part instances only (few samples, low occupancy: latency-bound): in the one-part
  // instances the ten extra live registers cost more than the hidden latency gains (the 1M-sample
  // launch hits L1 99.7 % of the time; measured 1327 -> 1361 us with deferral, 4.70e10 -> 4.95e10
  // waypoint-samples/s on config 5).
  constexpr bool DEFER = MP && GM != GM_NONE && !FKM;
  uint32_t pend[10];
#pragma unroll
  for (int l = 0; l < 10; ++l) pend[l] = raw_no_hit<GM == GM_NONE ? GM_U8_CDG : GM>();
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
    waypoint_cost2<D, GM, LIM, FKM, !DEFER>(P, x, xprev, vprev, vnew, &a2, &l2, &cw0, &cw1, &f0, &f1);
    if (DEFER) {
      uint32_t rl[5], rh[5];
      waypoint_lookups2<D, GM == GM_NONE ? GM_U8_CDG : GM>(P, x, rl, rh);
#pragma unroll
      for (int l = 0; l < 10; ++l) // the previous pair's lookups (cw0 / cw1 are zero here)
        cw0 += hit_of<GM == GM_NONE ? GM_U8_CDG : GM>(pend[l], P.coll_dist);
#pragma unroll
      for (int l = 0; l < 5; ++l) {
        pend[l] = rl[l];
        pend[5 + l] = has_hi ? rh[l] : raw_no_hit<GM == GM_NONE ? GM_U8_CDG : GM>();
      }
    }
    const bool valid0 = f0 & fine1 & fine2, valid1 = f1 & f0 & fine1;
    fine2 = f0 & (P.fast_ok != 0);
    fine1 = f1 & (P.fast_ok != 0);
    // an angle outside the ranges the straight-line arithmetic was verified for (never in
    // practice): remember it; the run is redone below with the general path.  Keeping the
    // out-of-line call out of this loop saves ~50 register moves per pair around the call site.
    bad = bad | !(valid0 & valid1);
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
  if (DEFER) { // the last pair's lookups
#pragma unroll
    for (int l = 0; l < 10; ++l) coll += hit_of<GM == GM_NONE ? GM_U8_CDG : GM>(pend[l], P.coll_dist);
  }
  if (bad) {
    // Rare path: the whole run again, every waypoint with IEEE divides and full-range sin / cos
    // (waypoint_cost_general).  Wherever the straight-line results are valid they equal these
    // bit for bit, so redoing all of the run gives what redoing the offending waypoints would.
    WpIn<D> in;
    if (!MP || w0 == 0) {
#pragma unroll
      for (int j = 0; j < D; ++j) in.xprev[j] = in.vprev[j] = __ldg(P.start + j);
    } else {
#pragma unroll
      for (int j = 0; j < D; ++j) {
        const float x1 = add_(sbase[j * NP + w0 - 1], eps[j]);
        const float x2 = add_(sbase[j * NP + w0 - 2], eps[j]);
        in.xprev[j] = x1;
        in.vprev[j] = div_(sub_(x1, x2), P.dt_div.b);
      }
    }
    total = 0.0f;
    ltotal = 0.0f;
    coll = 0;
#pragma unroll 1
    for (uint32_t w = w0; w < w_end; ++w) {
#pragma unroll
      for (int j = 0; j < D; ++j) in.x[j] = add_(sbase[j * NP + w], eps[j]);
      const WpOut<D> o = waypoint_cost_general<D, GM, true, LIM, FKM>(P, in);
      coll += o.coll;
      if (!MP) {
        total = add_(total, o.a2);
        if (LIM) ltotal = add_(ltotal, o.l2);
      } else {
        spart[((size_t)swi * WS + w) * 32 + lane] = o.a2;
        if (LIM) slim[((size_t)swi * WS + w) * 32 + lane] = o.l2;
      }
#pragma unroll
      for (int j = 0; j < D; ++j) {
        in.xprev[j] = in.x[j];
        in.vprev[j] = o.vnew[j];
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

} // namespace

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

cudaError_t DOAPP_DIM_FN(launch_cost_pk_dim)(const CostParams &p, int gm, cudaStream_t st) {
  return launch_cost_pk_d<DOAPP_D>(p, gm, st);
}

} // namespace doapp
