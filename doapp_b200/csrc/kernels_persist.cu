// kernels_persist.cu — plan_persistent_kernel: the whole keep-best loop in one cooperative
// launch (see kernels_cost.cu for the stage map).  Compiled once per joint count.
#include "cost_device.cuh"

#ifndef DOAPP_D
#error "compile with -DDOAPP_D=4 or -DDOAPP_D=5"
#endif

namespace doapp {

namespace {

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

} // namespace

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

const void *DOAPP_DIM_FN(persistent_kernel_dim)(int gm, bool lim, bool fkm) {
  return persistent_kernel_of<DOAPP_D>(gm, lim, fkm);
}

} // namespace doapp
