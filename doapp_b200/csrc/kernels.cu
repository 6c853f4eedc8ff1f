// kernels.cu — hand-written sm_100a kernels of the perturbator hot path.
//
// Stage map:
//   cost_kernel     a1 (Philox noise, fused) + a2 perturb + a3 derivatives +
//                   a4 FK + a5 grid gathers + a6 score + a7 per-climber argmin keys
//   update_kernel   a7 keep-best (generate_trajectories.cu:225-241, D5/D6 fixed)
//   final_kernel    host argmin of planning_state.cu:113-127 + velocities
//   scatter_kernel  a10 point cloud -> voxel occupancy (occupancy_grid.cpp:42-58)
//
// Work decomposition of cost_kernel: one warp owns one sample at a time; lane l
// owns waypoints l, l+32, l+64, ...  The noisy trajectory, velocities and
// accelerations the reference materialises in three global arrays
// (generate_trajectories.cu:202-210) live only in registers; neighbouring
// waypoints are exchanged with warp shuffles.  Sums over waypoints use a fixed
// order (lane-strided partials, then an xor butterfly) instead of the
// reference's unordered float atomicAdd (generate_trajectories.cu:73-77), so
// scores and the argmin are reproducible bit for bit.
#include "kernels.cuh"

#include <cstring>

#include "device_math.cuh"

namespace doapp {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kCostThreads = 256;

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

template <int D>
__device__ __forceinline__ void fk_points(const float (&th)[D], float rpy_s, float rpy_c,
                                          P3 (&out)[5]) {
  float s, c;
  P3 p, q;
  // base: t = (0, -0.0015, 0.0715), roty(theta0)            gt.cu:405-410
  dm::sincos(th[0], &s, &c);
  q = P3{0.0f, -0.0015f, 0.0715f};
  p = rot_y(s, c, q);
  out[0] = p;
  // shoulder: t = (0, 0, 0.045), rotz(theta1)                gt.cu:416-421
  dm::sincos(th[1], &s, &c);
  q = P3{p.x, p.y, dm::add_(p.z, 0.045f)};
  p = rot_z(s, c, q);
  out[1] = p;
  // elbow: t = (0, 0, 0.1), roty(theta2)                     gt.cu:427-432
  dm::sincos(th[2], &s, &c);
  q = P3{p.x, p.y, dm::add_(p.z, 0.1f)};
  p = rot_y(s, c, q);
  out[2] = p;
  // wrist: t = (0, 0, 0.1), roty(theta3)                     gt.cu:438-443
  dm::sincos(th[3], &s, &c);
  q = P3{p.x, p.y, dm::add_(p.z, 0.1f)};
  p = rot_y(s, c, q);
  out[3] = p;
  // wrist rotate: t = (0, -0.0125, 0.08), roty(theta4)       gt.cu:449-454
  q = P3{p.x, dm::add_(p.y, -0.0125f), dm::add_(p.z, 0.08f)};
  if (D >= 5) {
    dm::sincos(th[D >= 5 ? 4 : 0], &s, &c);
    p = rot_y(s, c, q);
  } else {
    p = q; // theta4 := 0 -> identity
  }
  // fixed Rrpy(0, -pi/2, 0) and t = (0, -0.005, 0.07)        gt.cu:456-459
  p = rot_y(rpy_s, rpy_c, p);
  p = P3{p.x, dm::add_(p.y, -0.005f), dm::add_(p.z, 0.07f)};
  out[4] = p;
}

template <int D>
__device__ __forceinline__ int num_collisions(const CostParams &P, const float (&x)[D]) {
  P3 pts[5];
  fk_points<D>(x, P.rpy_s, P.rpy_c, pts);
  int cnt = 0;
#pragma unroll
  for (int l = 0; l < 5; ++l) cnt += grid_collides(P.grid, pts[l].x, pts[l].y, pts[l].z, P.coll_dist);
  return cnt;
}

__device__ __forceinline__ float butterfly_sum(float v) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v = dm::add_(v, __shfl_xor_sync(kFull, v, off));
  return v;
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
template <int D>
__global__ void __launch_bounds__(kCostThreads)
    cost_kernel(const __grid_constant__ CostParams P) {
  const int lane = threadIdx.x & 31;
  const uint64_t warp = (uint64_t)blockIdx.x * (kCostThreads / 32) + (threadIdx.x >> 5);
  const uint64_t s_begin = warp * P.samples_per_warp;
  if (s_begin >= P.total_samples) return;
  uint64_t s_end = s_begin + P.samples_per_warp;
  if (s_end > P.total_samples) s_end = P.total_samples;

  float start[D];
#pragma unroll
  for (int j = 0; j < D; ++j) start[j] = __ldg(P.start + j);

  unsigned long long run_key = kKeyNone, run_keyf = kKeyNone;
  uint32_t run_c = 0xFFFFFFFFu;

  for (uint64_t s = s_begin; s < s_end; ++s) {
    const uint32_t c = (uint32_t)(s / P.m);
    const uint32_t t = (uint32_t)(s - (uint64_t)c * P.m);
    const uint32_t tg = (uint32_t)(P.sample_offset + t);
    if (c != run_c) {
      if (lane == 0 && run_c != 0xFFFFFFFFu) {
        if (P.keys) atomicMin(P.keys + run_c, run_key);
        if (P.keys_float) atomicMin(P.keys_float + run_c, run_keyf);
      }
      run_c = c;
      run_key = kKeyNone;
      run_keyf = kKeyNone;
    }
    const float *traj = P.traj + (size_t)c * D * P.npad;
    const float *nz = nullptr;
    if (P.noise_mode == NOISE_EXTERNAL)
      nz = P.noise + ((size_t)c * P.noise_m + (size_t)(tg - P.noise_t0)) * P.n_noise * D;

    float eps[D];
#pragma unroll
    for (int j = 0; j < D; ++j) eps[j] = 0.0f;
    if (P.n_noise == 1) {
      if (P.noise_mode == NOISE_EXTERNAL) {
#pragma unroll
        for (int j = 0; j < D; ++j) eps[j] = __ldg(nz + j);
      } else if (P.noise_mode == NOISE_PHILOX) {
#pragma unroll
        for (int j = 0; j < D; ++j)
          eps[j] = dm::noise_at(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, 0, j);
      }
    }

    float carry_x[D], carry_v[D];
#pragma unroll
    for (int j = 0; j < D; ++j) {
      carry_x[j] = start[j];
      carry_v[j] = start[j];
    }
    float part = 0.0f, lpart = 0.0f;
    int coll = 0;

    for (uint32_t w0 = 0; w0 < P.n; w0 += 32) {
      const uint32_t w = w0 + lane;
      const bool valid = w < P.n;
      float x[D];
      float a2 = 0.0f, l2 = 0.0f;
#pragma unroll
      for (int j = 0; j < D; ++j) {
        float b = valid ? __ldg(traj + (size_t)j * P.npad + w) : 0.0f;
        float e = eps[j];
        if (P.n_noise != 1 && valid) {
          if (P.noise_mode == NOISE_EXTERNAL)
            e = __ldg(nz + (size_t)w * D + j);
          else if (P.noise_mode == NOISE_PHILOX)
            e = dm::noise_at(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, w, j);
        }
        x[j] = dm::add_(b, e); // generate_trajectories.cu:53-57
        // prior position: lane-1, or the previous chunk's lane 31, or the
        // initial waypoint for waypoint 0 (generate_trajectories.cu:285)
        float xp = __shfl_up_sync(kFull, x[j], 1);
        float cx = __shfl_sync(kFull, carry_x[j], 31);
        if (lane == 0) xp = cx;
        float v = dm::div_(dm::sub_(x[j], xp), P.dt); // :288
        float vp = __shfl_up_sync(kFull, v, 1);
        float cv = __shfl_sync(kFull, carry_v[j], 31);
        if (lane == 0) vp = cv;
        float a = dm::div_(dm::sub_(v, vp), P.dt);
        a2 = dm::add_(a2, dm::mul_(a, a)); // :76
        carry_x[j] = x[j];
        carry_v[j] = v;
        if (P.use_limits) {
          float h = fmaxf(0.0f, fmaxf(dm::sub_(x[j], P.lim_max[j]), dm::sub_(P.lim_min[j], x[j])));
          l2 = dm::add_(l2, dm::mul_(h, h));
        }
      }
      if (valid) {
        part = dm::add_(part, a2);
        lpart = dm::add_(lpart, l2);
        if (P.collision_mode != 0 && P.grid.kind != 0) coll += num_collisions<D>(P, x);
      }
    }
    const float sum = butterfly_sum(part);
    const int coll_all = __reduce_add_sync(kFull, coll);
    float score = dm::add_((float)coll_all, dm::mul_(P.smooth_scale, sum)); // :81,100,107-108
    if (P.use_limits && P.limit_weight != 0.0f) {
      const float lsum = butterfly_sum(lpart);
      score = dm::add_(score, dm::mul_(P.limit_weight, lsum));
    }
    if (lane == 0) {
      if (P.scores) P.scores[s] = score;
      if (P.collisions) P.collisions[s] = coll_all;
    }
    unsigned long long key = ((unsigned long long)key_hi(P.select_mode, score) << 32) | tg;
    unsigned long long keyf = ((unsigned long long)key_hi_float(score) << 32) | tg;
    run_key = key < run_key ? key : run_key;
    run_keyf = keyf < run_keyf ? keyf : run_keyf;
  }
  if (lane == 0 && run_c != 0xFFFFFFFFu) {
    if (P.keys) atomicMin(P.keys + run_c, run_key);
    if (P.keys_float) atomicMin(P.keys_float + run_c, run_keyf);
  }
}

// -------------------------------------------------------------- update kernel
// One block per climber: keep-best of generate_trajectories.cu:225-241 (the
// winning sample replaces the climber's trajectory when its key beats the
// current one), then re-arm the key slot for the next iteration.
__global__ void update_kernel(const __grid_constant__ UpdateParams P) {
  const uint32_t c = blockIdx.x;
  __shared__ unsigned long long s_key;
  __shared__ uint32_t s_cur;
  if (threadIdx.x == 0) {
    s_key = P.keys[c];
    s_cur = P.cur_hi[c];
  }
  __syncthreads();
  const unsigned long long key = s_key;
  const uint32_t hi = (uint32_t)(key >> 32), tg = (uint32_t)key;
  if (P.init) {
    if (threadIdx.x == 0) {
      P.cur_hi[c] = hi;
      P.found_better[c] = 0;
      P.keys[c] = kKeyNone;
    }
    return;
  }
  const bool improve = hi < s_cur;
  if (improve) {
    float *traj = P.traj + (size_t)c * P.d * P.npad;
    const float *nz = nullptr;
    if (P.noise_mode == NOISE_EXTERNAL)
      nz = P.noise + ((size_t)c * P.noise_m + (size_t)(tg - P.noise_t0)) * P.n_noise * P.d;
    for (uint32_t i = threadIdx.x; i < P.n * P.d; i += blockDim.x) {
      const uint32_t j = i / P.n, w = i - j * P.n;
      const uint32_t wn = P.n_noise == 1 ? 0 : w;
      float e = 0.0f;
      if (P.noise_mode == NOISE_EXTERNAL)
        e = nz[(size_t)wn * P.d + j];
      else if (P.noise_mode == NOISE_PHILOX)
        e = dm::noise_at(P.seed_lo, P.seed_hi, P.sigma, P.iteration, c, tg, wn, j);
      traj[(size_t)j * P.npad + w] = dm::add_(traj[(size_t)j * P.npad + w], e);
    }
  }
  if (threadIdx.x == 0) {
    if (improve) {
      P.cur_hi[c] = hi;
      P.found_better[c] = 1;
    }
    P.keys[c] = kKeyNone;
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
// (occupancy_grid.cpp:47-58).  One thread per point, 12-byte coalesced reads,
// one byte store; equal values make concurrent stores to a voxel benign.
__global__ void scatter_kernel(const __grid_constant__ ScatterParams P) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P.num_points; i += stride) {
    float x = __ldg(P.xyz + 3 * i), y = __ldg(P.xyz + 3 * i + 1), z = __ldg(P.xyz + 3 * i + 2);
    long long idx = grid_index(P.grid, x, y, z);
    if (idx >= 0) P.cells[idx] = (uint8_t)P.val;
  }
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
__global__ void debug_fk_kernel(const float *wp, uint64_t count, float rpy_s, float rpy_c,
                                float *out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float th[D];
#pragma unroll
  for (int j = 0; j < D; ++j) th[j] = wp[i * D + j];
  P3 pts[5];
  fk_points<D>(th, rpy_s, rpy_c, pts);
  for (int l = 0; l < 5; ++l) {
    out[(i * 5 + l) * 3 + 0] = pts[l].x;
    out[(i * 5 + l) * 3 + 1] = pts[l].y;
    out[(i * 5 + l) * 3 + 2] = pts[l].z;
  }
}

inline unsigned blocks_for(uint64_t total, unsigned threads) {
  return (unsigned)((total + threads - 1) / threads);
}

} // namespace

void host_sincos(float x, float *s, float *c) { dm::sincos(x, s, c); }

cudaError_t launch_cost(const CostParams &p_in, int d, int sm_count, cudaStream_t st) {
  CostParams p = p_in;
  if (p.total_samples == 0) return cudaSuccess;
  // One warp per sample while that fits the resident warps of the chip, else
  // a contiguous run of samples per warp (blocked, so a warp's per-climber
  // running argmin needs one atomic per climber it touches).
  const uint64_t warps_per_block = kCostThreads / 32;
  const uint64_t resident_warps = (uint64_t)sm_count * 64; // 2048 threads / SM
  uint64_t spw = (p.total_samples + resident_warps - 1) / resident_warps;
  if (spw < 1) spw = 1;
  p.samples_per_warp = (uint32_t)spw;
  const uint64_t warps = (p.total_samples + spw - 1) / spw;
  const unsigned blocks = (unsigned)((warps + warps_per_block - 1) / warps_per_block);
  if (d == 5)
    cost_kernel<5><<<blocks, kCostThreads, 0, st>>>(p);
  else
    cost_kernel<4><<<blocks, kCostThreads, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_update(const UpdateParams &p, cudaStream_t st) {
  if (p.k == 0) return cudaSuccess;
  unsigned threads = p.n * p.d < 128 ? 64 : (p.n * p.d < 512 ? 128 : 256);
  update_kernel<<<p.k, threads, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_final(const FinalParams &p, cudaStream_t st) {
  final_kernel<<<1, 256, 0, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_scatter(const ScatterParams &p, int sm_count, cudaStream_t st) {
  if (p.num_points == 0) return cudaSuccess;
  unsigned blocks = blocks_for(p.num_points, 256);
  unsigned cap = (unsigned)sm_count * 8;
  if (blocks > cap) blocks = cap;
  scatter_kernel<<<blocks, 256, 0, st>>>(p);
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
