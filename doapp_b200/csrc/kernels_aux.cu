// kernels_aux.cu — launch geometry of the cost / planner kernels and the small kernels of the
// path: update_vector_kernel, final_kernel, scatter_kernel, depth_scatter_kernel, the distance
// transform, probes and parity helpers (see kernels_cost.cu for the stage map).
#include "cost_device.cuh"

namespace doapp {

namespace {

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

// ------------------------------------------------------ key exchange, sending side
// One thread per (climber, peer).  Order: re-arm own slots, fence, then the remote stores -- a
// peer can only write a re-armed slot after it has received the key sent here (kernels.cuh).
__global__ void xchg_send_kernel(const __grid_constant__ XchgSendParams P) {
  const uint32_t total = P.k * 8;
  for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) P.rearm_slot[i] = kKeyNone;
  if (P.rearm_set)
    for (uint32_t i = threadIdx.x; i < 3 * total; i += blockDim.x) P.rearm_set[i] = kKeyNone;
  __threadfence_system();
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < P.k * P.nranks; i += blockDim.x) {
    const uint32_t c = i / P.nranks, r = i - c * P.nranks;
    const unsigned long long key = P.keys[c];
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(P.peer_slot[r] + (size_t)c * 8 + P.rank), "l"(key) : "memory");
  }
  __threadfence_system();
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
  // 32-bit arithmetic throughout: (x - i)^2 < 2^20 (n <= kEdtMaxDim) and g <= kEdtInf = 2^29,
  // so every value below stays under 2^31 (64-bit multiplies and divides made this scan
  // several times slower)
  auto f = [&](int xx, int i) -> int { return (xx - i) * (xx - i) + g[i * 32 + lane]; };
  int q = 0;
  s[lane] = 0;
  t[lane] = 0;
  for (int u = 1; u < (int)n; ++u) {
    while (q >= 0 && f(t[q * 32 + lane], s[q * 32 + lane]) > f(t[q * 32 + lane], u)) --q;
    if (q < 0) {
      q = 0;
      s[lane] = (uint16_t)u;
    } else {
      const int i = s[q * 32 + lane];
      const int num = u * u - i * i + g[u * 32 + lane] - g[i * 32 + lane];
      const int den = 2 * (u - i);
      int w = num / den;
      if ((num - w * den != 0) && (num < 0)) --w; // floor division
      w += 1;
      if (w < (int)n) {
        ++q;
        s[q * 32 + lane] = (uint16_t)u;
        t[q * 32 + lane] = (uint16_t)w;
      }
    }
  }
  for (int u = (int)n - 1; u >= 0; --u) {
    const int v = f(u, s[q * 32 + lane]);
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

// Lines of up to kEdtBruteMaxN voxels: the lower envelope by exhaustive minimum instead of the
// sequential scan.  out[u] = min_i (u - i)^2 + g[i] = u^2 + min_i (h[i] - 2 u i) with
// h[i] = g[i] + i^2: ONE multiply-add and one minimum per (u, i) pair, no data-dependent control
// flow, no division, every lane busy -- n^2 pairs per line, which at n = 128 is far cheaper than
// a 128-step dependent scan whose every step waits on shared-memory round trips (measured: 76 +
// 110 us per 128^3 field with the scan, 45 + 48 us with this kernel in integer arithmetic).
// The arithmetic is fp32 and EXACT: finite squared distances are integers below 3 * 768^2 < 2^24,
// products 2 u i < 2^21, so every FFMA / FMNMX result is an integer fp32 represents exactly
// (scripts/micro/pipe_model.cu: IMAD occupies the FMA pipe for 2 cycles per warp instruction on
// sm_100a, FFMA for 1).  "No obstacle" (kEdtInf = 2^29) stays far above every finite value; its
// sums round, so anything >= 2^28 is reported as kEdtInf.  Same integers as the scan, hence the
// same bits.  A CTA owns 32 consecutive x (coalesced tile loads / stores, as the scan kernel) and
// the whole line length; warp w takes the lines w, w + 8, ...; lane l the outputs l, l + 32, ...
constexpr uint32_t kEdtBruteMaxN = 256;
template <bool TO_FLOAT>
__global__ void __launch_bounds__(256) edt_lines_bf_kernel(const int32_t *in, void *out_v, uint32_t nx,
                                                           uint32_t n, uint64_t stride,
                                                           uint64_t line_stride, float res) {
  extern __shared__ __align__(16) unsigned char edt_smem[];
  float *h = reinterpret_cast<float *>(edt_smem); // [n][33]
  const uint32_t x0 = blockIdx.x * 32, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t base = (uint64_t)blockIdx.y * line_stride + x0;
  for (uint32_t idx = threadIdx.x; idx < n * 32; idx += blockDim.x) {
    const uint32_t u = idx >> 5, xl = idx & 31;
    const int32_t g = x0 + xl < nx ? __ldg(in + base + xl + (uint64_t)u * stride) : kEdtInf;
    h[u * 33 + xl] = (float)(g + (int32_t)(u * u)); // exact below 2^24; kEdtInf + u^2 rounds (harmless)
  }
  __syncthreads();
  for (uint32_t xl = warp; xl < 32 && x0 + xl < nx; xl += 8) {
    float best[8], m2u[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      best[k] = 3.0e38f;
      m2u[k] = -2.0f * (float)(lane + 32 * k);
    }
    const float *col = h + xl;
    if (n <= 128) { // four outputs per lane
#pragma unroll 4
      for (uint32_t i = 0; i < n; ++i) {
        const float hv = col[i * 33], fi = (float)i;
#pragma unroll
        for (int k = 0; k < 4; ++k) best[k] = fminf(best[k], dm::fma_(m2u[k], fi, hv));
      }
    } else { // eight
#pragma unroll 2
      for (uint32_t i = 0; i < n; ++i) {
        const float hv = col[i * 33], fi = (float)i;
#pragma unroll
        for (int k = 0; k < 8; ++k) best[k] = fminf(best[k], dm::fma_(m2u[k], fi, hv));
      }
    }
    __syncwarp(); // every lane has read the whole column: it may now take the results
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t u = lane + 32 * k;
      if (u < n) {
        const float v = dm::add_(best[k], (float)(u * u));
        h[u * 33 + xl] = v >= 268435456.0f ? (float)kEdtInf : v;
      }
    }
  }
  __syncthreads();
  for (uint32_t idx = threadIdx.x; idx < n * 32; idx += blockDim.x) {
    const uint32_t u = idx >> 5, xl = idx & 31;
    if (x0 + xl >= nx) continue;
    const float v = h[u * 33 + xl];
    if (TO_FLOAT)
      reinterpret_cast<float *>(out_v)[base + xl + (uint64_t)u * stride] =
          v >= (float)kEdtInf ? INFINITY : dm::mul_(sqrtf(v), res);
    else
      reinterpret_cast<int32_t *>(out_v)[base + xl + (uint64_t)u * stride] = (int32_t)v;
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

} // namespace

void host_sincos(float x, float *s, float *c) { dm::sincos(x, s, c); }

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
  const void *kern = d == 5 ? persistent_kernel_dim5(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0)
                            : persistent_kernel_dim4(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0);
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
  const void *kern = d == 5 ? persistent_kernel_dim5(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0)
                            : persistent_kernel_dim4(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0);
  const uint32_t T = 1u << q.t_log2, SB = 256 / T;
  const size_t smem = persistent_smem(p, d, T, p.wpp);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  void *args[] = {(void *)&q};
  return cudaLaunchCooperativeKernel(kern, dim3((p.m + SB - 1) / SB, p.k), dim3(kPersistThreads), args, smem, st);
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
      return d == 5 ? launch_cost_pk_dim5(p, gm, st) : launch_cost_pk_dim4(p, gm, st);
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
  return d == 5 ? launch_cost_dim5(p, grid, threads, smem, cs, st)
                : launch_cost_dim4(p, grid, threads, smem, cs, st);
}

// ---- second-generation persistent planner: geometry, co-residency check, cooperative launch
static uint32_t pipeline_compute_warps(uint32_t sb, uint32_t tl) {
  const uint32_t T = 1u << tl;
  return T > 32 ? sb * (T >> 5) : (sb + (32u >> tl) - 1) / (32u >> tl);
}

static size_t pipeline_smem(const CostParams &p, int d, uint32_t sb, uint32_t tl) {
  const uint32_t NP = p.npad + 2, NPS = ((p.n + 3u) & ~3u) + 4u, T = 1u << tl, wps = T > 32 ? T >> 5 : 1u;
  const uint32_t cw = pipeline_compute_warps(sb, tl);
  size_t words = (size_t)2 * d * NP + (size_t)kP2Q * sb * NPS * (p.use_limits ? 2 : 1) +
                 (size_t)kP2Q * sb * 8 + 8 + (size_t)kP2Q * sb * wps + (size_t)kP2Q * cw + kP2Q + 8;
  return words * sizeof(float);
}

uint32_t pipeline_grid_x(const Persist2Params &q) { return (q.cost.m + q.sb - 1) / q.sb; }

cudaError_t pipeline_plan_config(const CostParams &p, int d, int sm_count, Persist2Params *q, bool *ok) {
  *ok = false;
  const int gm = grid_mode_of(p);
  // packed arithmetic only (as cost_kernel_pk): verified reciprocal divides, no generic grid;
  // the sample index shares its word with the 8-bit tag
  if (p.k == 0 || p.m == 0 || p.n_noise != 1 || gm == GM_GENERIC || !p.fast_ok) return cudaSuccess;
  if (p.m_global >= (1ull << 24) || p.sample_offset + p.m > (1ull << 24)) return cudaSuccess;
  if ((uint32_t)sm_count < p.k) return cudaSuccess;
  // one CTA per SM: the fewest samples per CTA whose grid still fits the chip
  const uint32_t ctas_per_climber = (uint32_t)sm_count / p.k;
  const uint32_t sb = (p.m + ctas_per_climber - 1) / ctas_per_climber;
  // one waypoint per lane, up to 64 lanes (two warps) per sample; fewer lanes (more waypoints
  // per lane) when the CTA's samples would need more than kP2MaxComputeWarps warps
  uint32_t tl = 0;
  while ((1u << tl) < p.n && tl < 6) ++tl;
  while (tl > 0 && pipeline_compute_warps(sb, tl) > (uint32_t)kP2MaxComputeWarps) --tl;
  const uint32_t cw = pipeline_compute_warps(sb, tl);
  if (cw > (uint32_t)kP2MaxComputeWarps) return cudaSuccess;
  const uint32_t T = 1u << tl, wpt = (p.n + T - 1) / T;
  const size_t smem = pipeline_smem(p, d, sb, tl);
  if (smem > 200 * 1024) return cudaSuccess;
  const void *kern = d == 5 ? pipeline_kernel_dim5(gm, p.use_limits != 0, p.fk_on != 0)
                            : pipeline_kernel_dim4(gm, p.use_limits != 0, p.fk_on != 0);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int threads = 32 * (int)(cw + kP2HelperWarps);
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem);
  if (e != cudaSuccess) return e;
  const uint64_t ctas = (uint64_t)p.k * ((p.m + sb - 1) / sb);
  if (per_sm < 1 || ctas > (uint64_t)per_sm * sm_count) return cudaSuccess;
  q->cost = p;
  q->cost.res_div = dm::FastDiv{p.grid.res, p.grid.res_rcp, 0.0f, 0.0f};
  q->t_log2 = tl;
  q->wpt = wpt;
  q->sb = sb;
  *ok = true;
  return cudaSuccess;
}

cudaError_t launch_pipeline_plan(const Persist2Params &q, int d, int sm_count, cudaStream_t st) {
  (void)sm_count;
  const CostParams &p = q.cost;
  const void *kern = d == 5 ? pipeline_kernel_dim5(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0)
                            : pipeline_kernel_dim4(grid_mode_of(p), p.use_limits != 0, p.fk_on != 0);
  const uint32_t cw = pipeline_compute_warps(q.sb, q.t_log2);
  const size_t smem = pipeline_smem(p, d, q.sb, q.t_log2);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  void *args[] = {(void *)&q};
  return cudaLaunchCooperativeKernel(kern, dim3(pipeline_grid_x(q), p.k), dim3(32 * (cw + kP2HelperWarps)),
                                     args, smem, st);
}

cudaError_t launch_verify_div(dm::FastDiv d, uint32_t lo_bits, uint32_t hi_bits, int mode,
                              unsigned long long *bad, int sm_count, cudaStream_t st) {
  verify_div_kernel<<<sm_count * 8, 256, 0, st>>>(d, lo_bits, hi_bits, mode, bad);
  return cudaGetLastError();
}

cudaError_t launch_xchg_send(const XchgSendParams &p, cudaStream_t st) {
  xchg_send_kernel<<<1, 256, 0, st>>>(p);
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
  static const bool scan_only = getenv("DOAPP_EDT_SCAN") != nullptr; // tuning / parity: the scan kernels only
  // along y: lines of plane z, element stride nx
  if (ny <= kEdtBruteMaxN && !scan_only)
    edt_lines_bf_kernel<false><<<dim3((nx + 31) / 32, nz), 256, (size_t)ny * 33 * 4, st>>>(
        tmp_a, tmp_b, nx, ny, nx, (uint64_t)nx * ny, res);
  else
    edt_lines_kernel<false><<<dim3((nx + 31) / 32, nz), 32, sy, st>>>(tmp_a, tmp_b, nx, ny, nx,
                                                                    (uint64_t)nx * ny, res);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  // along z: lines of row y, element stride nx * ny
  if (nz <= kEdtBruteMaxN && !scan_only)
    edt_lines_bf_kernel<true><<<dim3((nx + 31) / 32, ny), 256, (size_t)nz * 33 * 4, st>>>(
        tmp_b, out, nx, nz, (uint64_t)nx * ny, nx, res);
  else
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