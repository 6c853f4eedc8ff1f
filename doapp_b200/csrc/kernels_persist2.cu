// kernels_persist2.cu — plan_pipeline_kernel: the whole keep-best loop of
// generate_trajectories.cu:197-253 in ONE cooperative launch, second generation.
// Compiled once per joint count (-DDOAPP_D=4 / 5).
//
// Small problems (BASELINE config 2: 1024 samples x 64 waypoints x 100 iterations) are bound by
// latency, not throughput: one iteration is 65 536 waypoint-samples, under half a wave.  The
// kernel therefore runs ONE CTA PER SM with specialised warps and no CTA-wide barrier, grid
// barrier, contended atomic or fence anywhere in the loop (the only atomics are publishing
// stores into a slot nobody else writes):
//
//   compute warps   one WAYPOINT per lane (up to 64 lanes = two warps per sample), TWO
//                   ITERATIONS per pass: the halves of every packed fp32x2 register are the same
//                   waypoint under the noise of iteration e and of e + 1 (fma.rn.f32x2, the
//                   arithmetic of cost_kernel_pk); they cost the CURRENT trajectory
//                   speculatively, up to `depth` iterations ahead of the last decided one, and
//                   leave per-waypoint sums in a shared-memory ring;
//   fold warps (2)  lane s folds sample s's per-waypoint sums in the canonical order (left fold,
//                   waypoint 0 first: oracle ORC_SUM_WAYPOINT), takes the CTA's minimum key and
//                   publishes it with ONE 8-byte atomic exchange into the CTA's own slot of the
//                   iteration's entry ring in global memory;
//   collect warp    reads every CTA's slot of the oldest undecided iteration (coalesced 8-byte
//                   loads, no atomics) -- on several GPUs its own rank's peer mailbox instead,
//                   four iterations per poll -- applies the keep-best
//                   (generate_trajectories.cu:225-241, D5/D6 fixed) to the other trajectory
//                   buffer and bumps the version;
//   send warp       (several GPUs, first CTA) merges the rank's CTAs and stores the rank's key
//                   into every rank's mailbox over NVLink (one 8-byte reduction per peer),
//                   running ahead of the collect warp so that round trips overlap;
//   noise warps (3) draw the Philox / Box-Muller noise vectors of later iterations into a ring.
//
// Every published word carries its own validity: key = {score bits : 8-bit tag : 24-bit sample},
// tag = (iteration / ring) mod 16 and (trajectory version) mod 16.  A reader polls until all
// slots show the tag of the (iteration, version) it is deciding, so no slot is ever re-armed, no
// release/acquire pair is needed (the 8-byte word is the only data), and a CTA may abandon a
// speculative iteration the moment it learns that an earlier one was accepted: all CTAs (and
// ranks) see the same merged keys in the same order, so they agree on every (iteration, version)
// that matters; words published for abandoned versions are never waited for.
// An accepted update is rare (a few per plan), so the cost passes never wait for the exchange:
// with depth d the exchange may take d - 1 evaluation times.
// Arithmetic, noise, keys and tie-breaking are those of cost_kernel / cost_kernel_pk: results are
// bit-identical to the launch-per-iteration path (tests/test_plan_gpu.py).
#include "cost_device.cuh"

#ifndef DOAPP_D
#error "compile with -DDOAPP_D=4 or -DDOAPP_D=5"
#endif

namespace doapp {

namespace {

constexpr unsigned long long kTagMask = 0xFFull << 24;

__device__ __forceinline__ uint32_t p2_tag(uint32_t it, uint32_t ver) {
  return ((it / kP2Ring) & 15u) | ((ver & 15u) << 4);
}
__device__ __forceinline__ unsigned long long p2_pack(unsigned long long key, uint32_t tag) {
  return (key & ~kTagMask) | ((unsigned long long)tag << 24);
}
__device__ __forceinline__ unsigned long long p2_strip(unsigned long long e) { return e & ~kTagMask; }
__device__ __forceinline__ uint32_t p2_tag_of(unsigned long long e) { return (uint32_t)(e >> 24) & 0xFFu; }

__device__ __forceinline__ uint32_t ldv(const volatile uint32_t *p) { return *p; }
__device__ __forceinline__ void stv(volatile uint32_t *p, uint32_t v) { *p = v; }
__device__ __forceinline__ unsigned long long ld_gpu(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// publishing store: an atomic exchange is performed at the L2 at once, a plain store may sit in
// the SM's write path for microseconds (measured: 4 us from store to visibility)
__device__ __forceinline__ void st_gpu(unsigned long long *p, unsigned long long v) {
  unsigned long long old;
  asm volatile("atom.relaxed.gpu.global.exch.b64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// Remote publish without a return value: the slot is written by this sender only, so it knows
// the slot's content `cur` and turns it into `v` with ONE reduction (performed at the target, a
// peer's memory over NVLink).  An atomic exchange would work too, but its unused result
// serialises the sender on the NVLink round trip (3.3 us per iteration measured): the next
// exchange reuses the result register.
__device__ __forceinline__ void xor_sys(unsigned long long *p, unsigned long long cur, unsigned long long v) {
  asm volatile("red.relaxed.sys.global.xor.b64 [%0], %1;" ::"l"(p), "l"(cur ^ v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// control words of a CTA (shared memory, volatile)
struct P2State {
  uint32_t ver;     // trajectory version = accepted updates so far; buffer = ver & 1
  uint32_t base;    // first iteration to cost with this version (the one after the last accept)
  uint32_t decided; // iterations decided so far (every iteration below is final)
  uint32_t noise;   // iterations whose noise is in the ring
  uint32_t abort;   // exchange timed out: everybody leaves
  uint32_t pad[3];
};

// One cost pass of a compute lane: its waypoints w = r, r + T, ... of its sample for TWO
// iterations at once -- the two halves of every packed register are the same waypoint under the
// noise of iteration e (lo) and of iteration e + 1 (hi), both on the current trajectory (the
// second is one more step of speculation).  All arithmetic is "vertical" fma.rn.f32x2: no
// half-to-half moves.  The waypoint's priors (angles of w - 1, w - 2 and the velocity of w - 1;
// the initial waypoint for both derivative passes at w = 0, generate_trajectories.cu:285, D7)
// are recomputed from the shared trajectory, so lanes and warps are independent.
// Exactness: the straight-line divides / sin / cos are bit-identical to the IEEE formulation
// when every angle of the sample is "fine" (dm::fine_angle, device_math.cuh); the lane reports
// whether its own angles are, per half, and the fold warp redoes a sample that is not (never in
// practice) with the general path.
template <int D, int GM, bool LIM, bool FKM>
__device__ __forceinline__ void p2_evaluate(const CostParams &P, const float *sb, uint32_t NP,
                                            const dm::f2 (&eps)[D], const float (&start)[D],
                                            uint32_t r, uint32_t T, uint32_t wpt, float *sp_lo,
                                            float *sp_hi, float *sl_lo, float *sl_hi, int *coll_lo,
                                            int *coll_hi, bool *fine_lo, bool *fine_hi) {
  using namespace dm;
  int c_lo = 0, c_hi = 0;
  bool f_lo = true, f_hi = true;
  const uint32_t w_last = P.n - 1;
  for (uint32_t i = 0; i < wpt; ++i) {
    const uint32_t wq = r + i * T;
    const bool live = wq < P.n;
    const uint32_t w = live ? wq : w_last; // idle lanes shadow the last waypoint
    f2 x[D], a2 = f2_splat(0.0f), l2 = f2_splat(0.0f);
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const float *row = sb + j * NP;
      const f2 st = f2_splat(start[j]);
      x[j] = f2_add(f2_splat(row[w]), eps[j]); // generate_trajectories.cu:53-57
      const f2 x1 = w >= 1 ? f2_add(f2_splat(row[w - (w >= 1 ? 1 : 0)]), eps[j]) : st;
      const f2 x2 = w >= 2 ? f2_add(f2_splat(row[w - (w >= 2 ? 2 : 0)]), eps[j]) : st;
      const f2 v1 = w >= 1 ? f2_div_fast_unguarded(f2_sub(x1, x2), P.dt_div) : st;
      const f2 v = f2_div_fast_unguarded(f2_sub(x[j], x1), P.dt_div); // :288
      const f2 a = f2_div_fast_unguarded(f2_sub(v, v1), P.dt_div);
      a2 = f2_add(a2, f2_mul_rounded(a, a)); // :76 (product rounded, then added)
      f_lo = f_lo & fine_angle(f2_lo(x[j]));
      f_hi = f_hi & fine_angle(f2_hi(x[j]));
      if (LIM) {
        const f2 over = f2_sub(x[j], f2_splat(P.lim_max[j])), under = f2_sub(f2_splat(P.lim_min[j]), x[j]);
        const float h0 = fmaxf(0.0f, fmaxf(f2_lo(over), f2_lo(under)));
        const float h1 = fmaxf(0.0f, fmaxf(f2_hi(over), f2_hi(under)));
        l2 = f2_add(l2, f2_pack(__fmul_rn(h0, h0), __fmul_rn(h1, h1)));
      }
    }
    int cw0, cw1;
    waypoint_collisions2<D, GM, FKM>(P, x, &cw0, &cw1);
    if (live) {
      c_lo += cw0;
      c_hi += cw1;
      if (sp_lo) sp_lo[w] = f2_lo(a2);
      if (sp_hi) sp_hi[w] = f2_hi(a2);
      if (LIM && sl_lo) sl_lo[w] = f2_lo(l2);
      if (LIM && sl_hi) sl_hi[w] = f2_hi(l2);
    }
  }
  *coll_lo = c_lo;
  *coll_hi = c_hi;
  *fine_lo = f_lo;
  *fine_hi = f_hi;
}

template <int D, int GM, bool LIM, bool FKM>
__global__ void __launch_bounds__(kP2MaxThreads) plan_pipeline_kernel(const __grid_constant__ Persist2Params Q) {
  const CostParams &P = Q.cost;
  extern __shared__ __align__(16) float smem[];
  // T lanes per sample (1 .. 64), each lane the waypoints r, r + T, ...; WPS warps per sample
  // (T = 64) or spw samples per warp (T <= 32)
  const uint32_t T = 1u << Q.t_log2, WPS = T > 32 ? T >> 5 : 1u, spw = T > 32 ? 1u : 32u >> Q.t_log2;
  const uint32_t SB = Q.sb, CW = T > 32 ? SB * WPS : (SB + spw - 1) / spw; // samples / compute warps per CTA
  const uint32_t NP = P.npad + 2;                            // trajectory row
  const uint32_t NPS = ((P.n + 3u) & ~3u) + 4u;              // row of per-waypoint sums: 16-byte aligned
                                                             // for the fold's vector loads, the rows
                                                             // of 8 samples on distinct banks
  float *sbase = smem;                                       // [2][D][NP]   trajectory, by version & 1
  float *spart = sbase + 2 * D * NP;                         // [Q][SB][NPS] per-waypoint sum of acc^2
  float *slim = spart + (size_t)kP2Q * SB * NPS;             // [Q][SB][NPS] (limits)
  float *seps = slim + (LIM ? (size_t)kP2Q * SB * NPS : 0);  // [Q][SB][8]   noise of iteration it in it % Q
  float *swin = seps + (size_t)kP2Q * SB * 8;                // [8]          winner's noise
  int *scoll = (int *)(swin + 8);                            // [Q][SB][WPS] collision counts per compute warp
  volatile uint32_t *stag = (volatile uint32_t *)(scoll + kP2Q * SB * WPS); // [Q][CW] iteration << 8 | not-fine << 7 | version
  volatile uint32_t *snz = stag + kP2Q * CW;                 // [Q] iteration + 1 whose noise is in the slot
  volatile P2State *S = (volatile P2State *)(snz + kP2Q);

  const uint32_t c = blockIdx.y, k = gridDim.y, nctas = gridDim.x;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t iters = Q.iters, depth = Q.depth;

  // ---- prologue: the climber's trajectory -> buffer 0, control words
  for (uint32_t i = threadIdx.x; i < (uint32_t)D * NP; i += blockDim.x) {
    const uint32_t j = i / NP, w = i - j * NP;
    const uint32_t ws = w < P.n ? w : P.n - 1;
    sbase[i] = __ldg(P.traj + ((size_t)c * D + j) * P.npad + ws);
  }
  for (uint32_t i = threadIdx.x; i < kP2Q * CW; i += blockDim.x) stag[i] = 0xFFFFFFFFu;
  for (uint32_t i = threadIdx.x; i < kP2Q; i += blockDim.x) snz[i] = 0u;
  if (threadIdx.x == 0) {
    S->ver = 0;
    S->base = 0;
    S->decided = 0;
    S->noise = 0;
    S->abort = 0;
  }
  // the mailbox set of the NEXT launch: no peer writes it before this rank's last key of this
  // launch has left (it needs that key to finish its own launch)
  if (Q.nranks > 1 && blockIdx.x == 0)
    for (uint32_t i = threadIdx.x; i < kP2Ring * 8; i += blockDim.x)
      Q.mailbox_next[((size_t)(i >> 3) * k + c) * 8 + (i & 7)] = kKeyNone;
  __syncthreads(); // the only CTA-wide barrier of the kernel

  if (warp < CW) {
    // ================================================================ compute warps
    const uint32_t s_local = T > 32 ? warp / WPS : warp * spw + (lane >> Q.t_log2);
    const uint32_t wis = T > 32 ? warp % WPS : 0u;                    // warp index inside the sample
    const uint32_t r = T > 32 ? wis * 32 + lane : lane & (T - 1);     // lane index inside the sample
    const bool owns = s_local < SB;
    const uint32_t so = owns ? s_local : 0u;
    float start[D];
    bool start_fine = true;
#pragma unroll
    for (int j = 0; j < D; ++j) {
      start[j] = __ldg(P.start + j);
      start_fine = start_fine & dm::fine_angle(start[j]);
    }
    uint32_t my_ver = 0, e = 0, evals = 0;
    long long t_eval = 0, t_wait = 0, tc0 = clock64();
    for (;;) {
      bool leave = false, two = false;
      for (;;) { // wait for work: warp-uniform (every read is a broadcast of one shared word)
        const uint32_t v = ldv(&S->ver);
        if (v != my_ver) { // an earlier iteration was accepted: restart behind it
          __threadfence_block();
          e = ldv(&S->base);
          my_ver = v;
        }
        const uint32_t dec = ldv(&S->decided);
        if (dec >= iters || ldv(&S->abort)) {
          leave = true;
          break;
        }
        // a pass costs the same for one iteration as for two: unless only one is left (or the
        // depth is 1), wait until BOTH may run -- starting a single pass the moment the window
        // opens by one would keep the pipeline full of half-empty passes
        const bool pair = depth >= 2 && e + 1 < iters;
        if (e < iters && e + (pair ? 1u : 0u) < dec + depth && snz[e % kP2Q] == e + 1 &&
            (!pair || snz[(e + 1) % kP2Q] == e + 2)) {
          two = pair;
          break;
        }
        __nanosleep(32); // a spinning warp would take issue slots from the working ones
      }
      if (leave) break;
      __threadfence_block(); // the noise / trajectory words behind the control words
      const long long tc1 = clock64();
      t_wait += tc1 - tc0;
      const bool stamp = Q.stats && blockIdx.x == 0 && c == 0 && threadIdx.x == 0 && e < 256;
      if (stamp) Q.stats[8 + e * 8 + 0] = (unsigned long long)tc1; // pass of iteration e starts
      // iteration e in the lo halves, e + 1 in the hi halves (a copy of e when it may not run yet)
      const uint32_t q0 = e % kP2Q, q1 = (two ? e + 1 : e) % kP2Q;
      dm::f2 eps[D];
#pragma unroll
      for (int j = 0; j < D; ++j)
        eps[j] = dm::f2_pack(seps[((size_t)q0 * SB + so) * 8 + j], seps[((size_t)q1 * SB + so) * 8 + j]);
      int coll_lo, coll_hi;
      bool f_lo, f_hi;
      p2_evaluate<D, GM, LIM, FKM>(P, sbase + (my_ver & 1u) * D * NP, NP, eps, start, r, T, Q.wpt,
                                   owns ? spart + ((size_t)q0 * SB + s_local) * NPS : nullptr,
                                   owns && two ? spart + ((size_t)q1 * SB + s_local) * NPS : nullptr,
                                   owns ? slim + ((size_t)q0 * SB + s_local) * NPS : nullptr,
                                   owns && two ? slim + ((size_t)q1 * SB + s_local) * NPS : nullptr,
                                   &coll_lo, &coll_hi, &f_lo, &f_hi);
      f_lo = f_lo & start_fine;
      f_hi = f_hi & start_fine;
      if (T >= 32) {
        coll_lo = __reduce_add_sync(kFull, coll_lo);
        coll_hi = __reduce_add_sync(kFull, coll_hi);
      } else {
        for (uint32_t off = T >> 1; off >= 1; off >>= 1) {
          coll_lo += __shfl_xor_sync(kFull, coll_lo, off);
          coll_hi += __shfl_xor_sync(kFull, coll_hi, off);
        }
      }
      const bool bad_lo = !__all_sync(kFull, f_lo), bad_hi = !__all_sync(kFull, f_hi);
      if ((lane & (T - 1) & 31u) == 0 && owns) {
        scoll[(q0 * SB + s_local) * WPS + wis] = coll_lo;
        if (two) scoll[(q1 * SB + s_local) * WPS + wis] = coll_hi;
      }
      __syncwarp();
      __threadfence_block();
      if (lane == 0) {
        stag[q0 * CW + warp] = (e << 8) | (bad_lo ? 0x80u : 0u) | (my_ver & 0x7Fu);
        if (two) stag[q1 * CW + warp] = ((e + 1) << 8) | (bad_hi ? 0x80u : 0u) | (my_ver & 0x7Fu);
      }
      tc0 = clock64();
      if (stamp) Q.stats[8 + e * 8 + 1] = (unsigned long long)tc0; // ... and ends (tags written)
      e += two ? 2 : 1;
      ++evals;
      t_eval += tc0 - tc1;
    }
    if (Q.stats && blockIdx.x == 0 && c == 0 && threadIdx.x == 0) {
      Q.stats[1] = evals;
      Q.stats[2] = (unsigned long long)t_eval; // SM clocks inside cost passes / waiting (warp 0)
      Q.stats[3] = (unsigned long long)t_wait;
    }
  } else if (warp < CW + kP2FoldWarps) {
    // ================================================================ fold warps
    // (iterations alternate between them: fold warp f takes e = f mod kP2FoldWarps)
    const uint32_t fw = warp - CW;
    float start[D];
#pragma unroll
    for (int j = 0; j < D; ++j) start[j] = __ldg(P.start + j);
    uint32_t my_ver = 0, e = fw;
    long long t_fold = 0, t_fwait = 0, tf0 = clock64();
    for (;;) {
      bool leave = false;
      uint32_t bad = 0; // compute warps that met an angle outside the verified ranges
      for (;;) {
        const uint32_t v = ldv(&S->ver);
        if (v != my_ver) {
          __threadfence_block();
          const uint32_t b0 = ldv(&S->base);
          e = b0 + (fw + kP2FoldWarps - b0 % kP2FoldWarps) % kP2FoldWarps;
          my_ver = v;
        }
        if (ldv(&S->decided) >= iters || ldv(&S->abort)) {
          leave = true;
          break;
        }
        if (e < iters) {
          const uint32_t want = (e << 8) | (my_ver & 0x7Fu);
          bool ok = true, b = false;
          for (uint32_t wi = lane; wi < CW; wi += 32) {
            const uint32_t tg = stag[(e % kP2Q) * CW + wi];
            ok = ok && (tg & ~0x80u) == want;
            b = b || (tg & 0x80u) != 0;
          }
          if (__all_sync(kFull, ok)) {
            bad = __ballot_sync(kFull, b);
            break;
          }
        }
        __nanosleep(40);
      }
      if (leave) break;
      __threadfence_block();
      const long long tf1 = clock64();
      t_fwait += tf1 - tf0;
      const bool fstamp = Q.stats && blockIdx.x == 0 && c == 0 && lane == 0 && e < 256;
      if (fstamp) Q.stats[8 + e * 8 + 2] = (unsigned long long)tf1; // fold sees every warp's tag
      const uint32_t q = e % kP2Q;
      unsigned long long key = kKeyNone;
      for (uint32_t s = lane; s < SB; s += 32) {
        const uint32_t t = blockIdx.x * SB + s;
        const uint32_t tg = (uint32_t)(P.sample_offset + (t < P.m ? t : P.m - 1));
        const float *row = spart + ((size_t)q * SB + s) * NPS, *lrow = slim + ((size_t)q * SB + s) * NPS;
        float total = 0.0f, ltotal = 0.0f; // fixed order: left fold over the waypoints
        int coll = 0;
        if (bad == 0) {
          uint32_t w = 0;
          for (; w + 32 <= P.n; w += 32) { // loads first: the dependent chain is the adds only
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const float4 *>(row + w + 4 * u);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              total = dm::add_(total, v[u].x);
              total = dm::add_(total, v[u].y);
              total = dm::add_(total, v[u].z);
              total = dm::add_(total, v[u].w);
            }
            if (LIM) {
#pragma unroll
              for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const float4 *>(lrow + w + 4 * u);
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                ltotal = dm::add_(ltotal, v[u].x);
                ltotal = dm::add_(ltotal, v[u].y);
                ltotal = dm::add_(ltotal, v[u].z);
                ltotal = dm::add_(ltotal, v[u].w);
              }
            }
          }
          for (; w < P.n; ++w) {
            total = dm::add_(total, row[w]);
            if (LIM) ltotal = dm::add_(ltotal, lrow[w]);
          }
          for (uint32_t u = 0; u < WPS; ++u) coll += scoll[(q * SB + s) * WPS + u];
        } else {
          // rare: some angle of this CTA's samples is outside the ranges the straight-line
          // arithmetic was verified for: redo the sample with IEEE divides and full-range sin / cos
          const float *sb = sbase + (my_ver & 1u) * D * NP;
          WpIn<D> in;
#pragma unroll
          for (int j = 0; j < D; ++j) in.xprev[j] = in.vprev[j] = start[j];
          for (uint32_t w = 0; w < P.n; ++w) {
#pragma unroll
            for (int j = 0; j < D; ++j) in.x[j] = dm::add_(sb[j * NP + w], seps[((size_t)q * SB + s) * 8 + j]);
            const WpOut<D> o = waypoint_cost_general<D, GM, true, LIM, FKM>(P, in);
            total = dm::add_(total, o.a2);
            if (LIM) ltotal = dm::add_(ltotal, o.l2);
            coll += o.coll;
#pragma unroll
            for (int j = 0; j < D; ++j) {
              in.xprev[j] = in.x[j];
              in.vprev[j] = o.vnew[j];
            }
          }
        }
        float score = dm::add_((float)coll, dm::mul_(P.smooth_scale, total)); // :81,100,107-108
        if (LIM && P.limit_weight != 0.0f) score = dm::add_(score, dm::mul_(P.limit_weight, ltotal));
        if (t < P.m) {
          const unsigned long long kk = ((unsigned long long)key_hi(P.select_mode, score) << 32) | tg;
          key = kk < key ? kk : key;
        }
      }
      key = warp_min_u64(key);
      if (lane == 0)
        st_gpu(Q.entries + ((size_t)(e % kP2Ring) * k + c) * nctas + blockIdx.x, p2_pack(key, p2_tag(e, my_ver)));
      tf0 = clock64();
      if (fstamp) Q.stats[8 + e * 8 + 3] = (unsigned long long)tf0; // key published
      e += kP2FoldWarps;
      t_fold += tf0 - tf1;
    }
    if (Q.stats && blockIdx.x == 0 && c == 0 && lane == 0 && fw == 0) {
      Q.stats[4] = (unsigned long long)t_fold;
      Q.stats[5] = (unsigned long long)t_fwait;
    }
  } else if (warp == CW + kP2FoldWarps) {
    // ================================================================ collect warp
    uint32_t cur_hi = (uint32_t)(__ldg(P.pend_keys + c) >> 32); // the seed pass (gt.cu:180-191)
    uint32_t ver = 0, accepts = 0;
    bool dead = false;
    const unsigned long long t_start = globaltimer_ns();
    long long t_poll = 0, t_decide = 0, tk0 = clock64();
    uint32_t i = 0;
    while (i < iters && !dead) {
      // ---- the merged keys of iteration i (and, on several ranks, of up to three more)
      unsigned long long keys4[4] = {kKeyNone, kKeyNone, kKeyNone, kKeyNone};
      uint32_t ready = 0;
      if (Q.nranks <= 1) {
        const uint32_t want = p2_tag(i, ver);
        const unsigned long long *row = Q.entries + ((size_t)(i % kP2Ring) * k + c) * nctas;
        unsigned long long mn = kKeyNone;
        for (uint32_t spins = 1;; ++spins) { // every CTA's key of (i, ver): coalesced polls, no atomics
          bool ok = true;
          mn = kKeyNone;
          for (uint32_t u0 = 0; u0 < nctas; u0 += 32 * 8) { // up to 8 loads in flight per lane
            unsigned long long v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const uint32_t idx = u0 + u * 32 + lane;
              v[u] = idx < nctas ? ld_gpu(row + idx) : p2_pack(kKeyNone, want);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              ok = ok && p2_tag_of(v[u]) == want;
              const unsigned long long sv = p2_strip(v[u]);
              mn = sv < mn ? sv : mn;
            }
          }
          if (__all_sync(kFull, ok)) break;
          if ((spins & 0xFFFu) == 0 &&
              (*(volatile unsigned int *)Q.error != 0 || globaltimer_ns() - t_start > Q.timeout_ns)) {
            dead = true;
            break;
          }
          __nanosleep(Q.poll_ns); // a spinning warp takes issue slots from the compute warps of its scheduler
        }
        if (dead) break;
        keys4[0] = warp_min_u64(mn);
        ready = 1;
      } else {
        // several ranks: the send warp of the climber's first CTA has merged this rank's CTAs
        // and stored the rank's key into every rank's mailbox (the own one included); every CTA
        // polls its own rank's mailbox -- local memory.  One round looks at FOUR iterations
        // (lane group g = lane / 8: iteration i + g, one lane per rank), so the collect warp is
        // not limited to one iteration per poll round trip.
        const uint32_t g = lane >> 3, rl = lane & 7u;
        for (uint32_t spins = 1;; ++spins) {
          const uint32_t it = i + g;
          const uint32_t want = p2_tag(it, ver);
          unsigned long long v = p2_pack(kKeyNone, want);
          if (it < iters && rl < Q.nranks)
            v = ld_sys(Q.mailbox[Q.rank] + ((size_t)(it % kP2Ring) * k + c) * 8 + rl);
          const uint32_t okb = __ballot_sync(kFull, it < iters && p2_tag_of(v) == want);
          unsigned long long mn = p2_strip(v);
#pragma unroll
          for (int off = 4; off >= 1; off >>= 1) { // minimum inside the group of 8 lanes
            const unsigned long long o = __shfl_xor_sync(kFull, mn, off);
            mn = o < mn ? o : mn;
          }
          while (ready < 4 && ((okb >> (8 * ready)) & 0xFFu) == 0xFFu) ++ready; // consecutive ones
          if (ready) {
#pragma unroll
            for (int q = 0; q < 4; ++q) keys4[q] = __shfl_sync(kFull, mn, 8 * q);
            break;
          }
          if ((spins & 0xFFFu) == 0 &&
              (*(volatile unsigned int *)Q.error != 0 || globaltimer_ns() - t_start > Q.timeout_ns)) {
            dead = true;
            break;
          }
          __nanosleep(Q.poll_ns);
        }
        if (dead) break;
      }
      const long long tk1 = clock64();
      t_poll += tk1 - tk0;
      // ---- decide them in order; an accepted update voids the later ones (their version)
      for (uint32_t q = 0; q < ready; ++q, ++i) {
        const unsigned long long mn = keys4[q];
        const bool kstamp = Q.stats && blockIdx.x == 0 && c == 0 && lane == 0 && i < 256;
        if (kstamp) Q.stats[8 + i * 8 + 4] = (unsigned long long)tk1; // every key seen
        const uint32_t hi = (uint32_t)(mn >> 32);
        const bool accept = hi < cur_hi; // keep-best (generate_trajectories.cu:225-241, D5/D6 fixed)
        if (accept) {
          cur_hi = hi;
          ++accepts;
          const uint32_t win = (uint32_t)mn & 0xFFFFFFu;
          if (lane < 3) {
            float a = 0.0f, b = 0.0f;
            if (P.noise_mode == NOISE_PHILOX) {
              dm::noise_pair(P.seed_lo, P.seed_hi, P.sigma, i, c, win, 0, lane, &a, &b);
            } else if (P.noise_mode == NOISE_EXTERNAL) {
              const float *pn = Q.noise_base + (size_t)i * Q.noise_iter_stride +
                                ((size_t)c * P.noise_m + (size_t)(win - P.noise_t0)) * D;
              a = 2 * lane < D ? __ldg(pn + 2 * lane) : 0.0f;
              b = 2 * lane + 1 < D ? __ldg(pn + 2 * lane + 1) : 0.0f;
            }
            if (2 * lane < D) swin[2 * lane] = a;
            if (2 * lane + 1 < D) swin[2 * lane + 1] = b;
          }
          __syncwarp();
          const float *src = sbase + (ver & 1u) * D * NP;
          float *dst = sbase + ((ver + 1) & 1u) * D * NP;
          for (uint32_t idx = lane; idx < (uint32_t)D * NP; idx += 32) dst[idx] = dm::add_(src[idx], swin[idx / NP]);
          ++ver;
          __syncwarp();
          __threadfence_block();
          if (lane == 0) {
            S->base = i + 1;
            __threadfence_block();
            S->ver = ver;
            S->decided = i + 1;
          }
        } else if (lane == 0) {
          S->decided = i + 1;
        }
        if (kstamp) Q.stats[8 + i * 8 + 5] = (unsigned long long)clock64(); // decided
        if (accept) { // the keys fetched for later iterations belong to the old version
          ++i;
          break;
        }
      }
      tk0 = clock64();
      t_decide += tk0 - tk1;
    }
    if (Q.stats && blockIdx.x == 0 && c == 0 && lane == 0) {
      Q.stats[6] = (unsigned long long)t_poll;
      Q.stats[7] = (unsigned long long)t_decide;
    }
    if (dead) {
      if (lane == 0) {
        atomicExch(Q.error, 1u);
        S->abort = 1;
      }
    } else if (blockIdx.x == 0) { // the climber's final state, for the final scoring pass
      const float *fin = sbase + (ver & 1u) * D * NP;
      for (uint32_t idx = lane; idx < (uint32_t)D * P.npad; idx += 32) {
        const uint32_t j = idx / P.npad, w = idx - j * P.npad;
        P.traj_next[(size_t)c * D * P.npad + idx] = w < P.n ? fin[j * NP + w] : 0.0f;
      }
      if (lane == 0) {
        P.cur_hi_out[c] = cur_hi;
        P.found_better[c] = accepts ? 1 : 0;
        if (Q.stats && c == 0) Q.stats[0] = accepts;
      }
    }
  } else if (warp == CW + kP2FoldWarps + 1) {
    // ================================================================ send warp (several ranks)
    // Runs ahead of the collect warp: as soon as every CTA of this rank has published its key of
    // (a, ver), the rank's key goes to every rank's mailbox -- one 8-byte word per peer over
    // NVLink, carrying its own tag -- without waiting for the answers to earlier iterations, so
    // the NVLink round trips of successive iterations overlap.
    if (Q.nranks > 1 && blockIdx.x == 0) {
      __shared__ unsigned long long slast[kP2Ring]; // what this rank's slot of ring index r holds
      for (uint32_t r = lane; r < (uint32_t)kP2Ring; r += 32) slast[r] = kKeyNone; // as armed
      __syncwarp();
      uint32_t my_ver = 0, a = 0;
      for (;;) {
        const uint32_t v = ldv(&S->ver);
        if (v != my_ver) {
          __threadfence_block();
          a = ldv(&S->base);
          my_ver = v;
        }
        if (ldv(&S->decided) >= iters || ldv(&S->abort)) break;
        if (a >= iters) {
          __nanosleep(200);
          continue;
        }
        const uint32_t want = p2_tag(a, my_ver);
        const unsigned long long *row = Q.entries + ((size_t)(a % kP2Ring) * k + c) * nctas;
        bool ok = true;
        unsigned long long mn = kKeyNone;
        for (uint32_t u0 = 0; u0 < nctas; u0 += 32 * 8) {
          unsigned long long w8[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t idx = u0 + u * 32 + lane;
            w8[u] = idx < nctas ? ld_gpu(row + idx) : p2_pack(kKeyNone, want);
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            ok = ok && p2_tag_of(w8[u]) == want;
            const unsigned long long sv = p2_strip(w8[u]);
            mn = sv < mn ? sv : mn;
          }
        }
        if (!__all_sync(kFull, ok)) {
          __nanosleep(Q.poll_ns);
          continue;
        }
        mn = warp_min_u64(mn);
        const unsigned long long word = p2_pack(mn, want), cur = slast[a % kP2Ring];
        if (lane < Q.nranks)
          xor_sys(Q.mailbox[lane] + ((size_t)(a % kP2Ring) * k + c) * 8 + Q.rank, cur, word);
        __syncwarp();
        if (lane == 0) slast[a % kP2Ring] = word;
        __syncwarp();
        ++a;
      }
    }
  } else {
    // ================================================================ noise warp
    // one d-vector of noise per sample and iteration (generate_trajectories.cu:37-44), drawn up
    // to Q - 1 iterations ahead of the last decided one: one Box-Muller pair per lane
    const uint32_t nw = warp - (CW + kP2FoldWarps + 2);
    for (uint32_t g = nw; g < iters; g += kP2NoiseWarps) {
      bool leave = false;
      while (g >= ldv(&S->decided) + kP2Q) {
        if (ldv(&S->abort)) {
          leave = true;
          break;
        }
        __nanosleep(200);
      }
      if (leave) break;
      float *se = seps + (size_t)(g % kP2Q) * SB * 8;
      if (P.noise_mode == NOISE_PHILOX) {
        for (uint32_t l = lane; l < SB * 3; l += 32) {
          const uint32_t s = l / 3, pr = l - 3 * s;
          uint32_t t = blockIdx.x * SB + s;
          t = t < P.m ? t : P.m - 1;
          float a, b;
          dm::noise_pair(P.seed_lo, P.seed_hi, P.sigma, g, c, (uint32_t)(P.sample_offset + t), 0, pr, &a, &b);
          if (2 * pr < D) se[s * 8 + 2 * pr] = a;
          if (2 * pr + 1 < D) se[s * 8 + 2 * pr + 1] = b;
        }
      } else {
        for (uint32_t l = lane; l < SB * D; l += 32) {
          const uint32_t s = l / D, j = l - D * s;
          uint32_t t = blockIdx.x * SB + s;
          t = t < P.m ? t : P.m - 1;
          se[s * 8 + j] = P.noise_mode == NOISE_EXTERNAL
                              ? __ldg(Q.noise_base + (size_t)g * Q.noise_iter_stride +
                                      ((size_t)c * P.noise_m + (size_t)(P.sample_offset + t - P.noise_t0)) * D + j)
                              : 0.0f;
        }
      }
      __syncwarp();
      __threadfence_block();
      if (lane == 0) snz[g % kP2Q] = g + 1;
    }
  }
}

} // namespace

template <int D> static const void *pipeline_kernel_of(int gm, bool lim, bool fkm) {
#define DOAPP_P2(GMV)                                                  \
  return lim ? (const void *)plan_pipeline_kernel<D, GMV, true, false> \
             : (const void *)plan_pipeline_kernel<D, GMV, false, false>
#define DOAPP_P2_FK(GMV)                                              \
  return lim ? (const void *)plan_pipeline_kernel<D, GMV, true, true> \
             : (const void *)plan_pipeline_kernel<D, GMV, false, true>
  if (fkm) {
    if (gm == GM_U8_CDG) DOAPP_P2_FK(GM_U8_CDG);
    if (gm == GM_F32_CDG) DOAPP_P2_FK(GM_F32_CDG);
  }
  switch (gm) {
  case GM_U8_CDG: DOAPP_P2(GM_U8_CDG);
  case GM_F32_CDG: DOAPP_P2(GM_F32_CDG);
  default: DOAPP_P2(GM_NONE);
  }
#undef DOAPP_P2
#undef DOAPP_P2_FK
}

const void *DOAPP_DIM_FN(pipeline_kernel_dim)(int gm, bool lim, bool fkm) {
  return pipeline_kernel_of<DOAPP_D>(gm, lim, fkm);
}

} // namespace doapp
