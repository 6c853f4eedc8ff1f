// doapp_capi.cu — implementation of the C ABI declared in include/doapp/doapp.h.
//
// Host side of the hot path: owns every device buffer (plain cudaMalloc; the
// reference uses cudaMallocManaged for everything, vector.cuh:142-167), stages
// inputs, enqueues the kernels of kernels_*.cu on one stream per ctx — the whole
// planner loop as a single CUDA graph — and reads results back.  There is no
// CPU fallback: without a device doapp_create fails.
#include "doapp/doapp.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include <cuda_runtime.h>
#include <nccl.h>

#include "kernels.cuh"

using namespace doapp;

namespace {

thread_local std::string g_create_error;

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t,
                            cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi *nccl_api(std::string *err) {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    // torch's bundled libnccl.so.2 is reused when it is already mapped
    api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!api.handle) api.handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (api.handle) {
      api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.handle, "ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.handle, "ncclCommInitRank");
      api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
      api.AllReduce = (decltype(api.AllReduce))dlsym(api.handle, "ncclAllReduce");
      api.Broadcast = (decltype(api.Broadcast))dlsym(api.handle, "ncclBroadcast");
      api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
    }
  }
  if (!api.handle || !api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.Broadcast) {
    if (err) *err = "libnccl.so.2 not loadable";
    return nullptr;
  }
  return &api;
}

template <typename T> struct DevBuf {
  T *p = nullptr;
  size_t cap = 0; // elements
};

} // namespace

struct doapp_ctx {
  doapp_config cfg{};
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  uint64_t launches = 0;

  uint32_t npad = 0;
  uint64_t S = 0; // local samples per iteration
  uint64_t m_global = 0;

  DevBuf<float> traj, aos_tmp, start, noise, scores, final_scores, out, cloud, pts;
  DevBuf<int32_t> coll, final_coll;
  DevBuf<unsigned long long> keys, keys_float;
  DevBuf<uint32_t> cur_hi;
  DevBuf<unsigned long long> pkeys; // persistent planner: 3 x [k] rotating barrier words {key, arrivals - 1}
  int persist_mode = -1;            // DOAPP_PERSIST=0/1: never / whenever legal
  int persist_version = 2;          // DOAPP_PERSIST_V=1: first-generation kernel (plan_persistent_kernel)
  int persist_depth = 0;            // DOAPP_PERSIST_DEPTH=1..15 (0: 8 on one rank, 10 on several)
  DevBuf<unsigned long long> p2_entries, p2_stats; // plan_pipeline_kernel: tagged CTA keys, counters
  DevBuf<unsigned int> p2_error;    // set by the kernel when the key exchange timed out
  bool p2_ran = false;              // the last plan ran plan_pipeline_kernel
  bool persist_graph_ok = true;     // the planner's launch sequence can be replayed as a CUDA graph
  // peer-memory argmin merge: own mailbox [2 epochs][3][k][8] + result [3][k], the ranks' mailboxes
  DevBuf<unsigned long long> mailbox, presult;
  unsigned long long *peer_mbox[8] = {};
  bool peer_ready = false;
  int peer_rank = 0, peer_nranks = 1;
  uint64_t peer_epoch = 0; // persistent multi-rank launches so far (same on every rank)
  DevBuf<long long> usum; // weighted update: 2 x [k][8] integer sums (ping-pong)
  // Ping-pong state of the planner: traj holds two [k][d][npad] halves, cur_hi two [k]
  // halves, keys three [k] buffers (read / write / being re-armed).  Every launch that
  // applies a pending keep-best reads half traj_s and writes the other (kernels.cuh).
  int traj_s = 0, cur_s = 0, key_w = -1;
  size_t traj_half = 0; // elements per trajectory half
  DevBuf<uint8_t> found, grid_cells, lookup_coll;
  DevBuf<uint16_t> depth; // staged depth frames (doapp_grid_from_depth)
  DevBuf<long long> lookup_idx;
  DevBuf<float> dbg_a, dbg_b, dbg_c;

  GridDev grid{};
  size_t grid_bytes = 0;
  uint64_t grid_generation = 0;

  float lim_min[kMaxJoints]{}, lim_max[kMaxJoints]{};
  int use_limits = 0;
  float rpy_s = 0.0f, rpy_c = 0.0f;
  DevBuf<FkModelDev> fk_dev;       // doapp_set_fk_model (URDF chain + spheres); fk_on selects it
  FkModelDev fk_host{};            // the same, for the kernels' parameter block
  bool fk_on = false;
  uint32_t fk_spheres = 0;
  dm::FastDiv dt_div{};            // verified at create (verify_fast_div)
  int pk_mode = -1;                // DOAPP_PK=0/1 forces the scalar / packed cost kernel
  bool dt_fast_ok = false;         // ... and delta_t inside [kFastDtLo, kFastDtHi]
  DevBuf<unsigned long long> vbad; // mismatch counter of the verification kernels
  bool fast_math_ok = true;        // DOAPP_EXACT_DIV=1 forces the IEEE divides

  // staged planner inputs
  bool plan_staged = false, plan_pending = false;
  int plan_noise_mode = NOISE_PHILOX;
  uint32_t plan_iters = 0;

  // cached CUDA graph of the planner loop
  cudaGraphExec_t graph_exec = nullptr;
  uint32_t graph_iters = 0;
  int graph_noise_mode = -1;
  const float *graph_noise_ptr = nullptr;
  uint64_t graph_grid_generation = ~0ull;
  uint64_t graph_kernels = 0;
  int graph_end_traj_s = 0, graph_end_cur_s = 0, graph_end_key_w = -1;
  // fused multi-GPU key exchange (xchg_send_kernel + mailbox polls in the cost launches): plans
  // alternate between two mailbox sets, so two graphs are cached; the one not current lives here
  struct GraphAlt {
    cudaGraphExec_t exec = nullptr;
    uint32_t iters = 0;
    int noise_mode = -1;
    const float *noise_ptr = nullptr;
    uint64_t grid_generation = ~0ull, kernels = 0;
    int end_traj_s = 0, end_cur_s = 0, end_key_w = -1;
  } graph_alt;
  int graph_parity = 0;   // mailbox set the current graph was captured for
  int xr_parity = 0;      // mailbox set of the plan being enqueued
  bool xr_on = false;     // the plan being enqueued merges its keys through the peer mailboxes
  int xr_mode = -1;       // DOAPP_XCHG=0: always NCCL
  DevBuf<unsigned int> xr_error;
  bool use_graph = true;
  // cached graph of the dynamic-obstacle replan loop (doapp_replan_dynamic)
  cudaGraphExec_t replan_exec = nullptr;
  uint32_t replan_iters = 0;
  size_t replan_points = 0;
  const float *replan_cloud = nullptr;
  const uint8_t *replan_cells = nullptr;
  uint64_t replan_grid_generation = ~0ull;
  uint64_t replan_kernels = 0;
  int replan_end[3] = {0, 0, -1};
  // doapp_replan_dynamic streams the clouds in on a second stream, one event per cloud, so the
  // H2D copy of cloud i + 1 overlaps the rebuild + replan of cloud i
  cudaStream_t copy_stream = nullptr;
  std::vector<cudaEvent_t> cloud_events;
  size_t cloud_staged = 0;        // elements of the clouds now in ctx->cloud
  bool replan_ext_wait_ok = true; // external event waits can be captured into the replan graph
  bool replan_waits = false;      // the cached replan graph carries the per-cloud waits

  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1;

  // per-launch CUDA-event timing of the cost kernel (doapp_plan_profile)
  std::vector<cudaEvent_t> prof_events;
  bool profiling = false;
  uint32_t last_plan_mode = 0; // doapp_plan_mode
  uint32_t prof_launches = 0; // cost launches the last profiled plan timed

  std::vector<float> h_out;
};

namespace {

doapp_status fail(doapp_ctx *ctx, doapp_status st, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx)
    ctx->err = buf;
  else
    g_create_error = buf;
  return st;
}

#define CU(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(ctx, e_ == cudaErrorMemoryAllocation ? DOAPP_OOM : DOAPP_CUDA_ERROR,            \
                  "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));             \
  } while (0)

#define ST(call)                                                                                  \
  do {                                                                                            \
    doapp_status s_ = (call);                                                                     \
    if (s_ != DOAPP_OK) return s_;                                                                \
  } while (0)

template <typename T> doapp_status ensure(doapp_ctx *ctx, DevBuf<T> &b, size_t elems) {
  if (elems <= b.cap && b.p) return DOAPP_OK;
  if (b.p) {
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
  }
  if (elems == 0) elems = 1;
  CU(cudaMalloc((void **)&b.p, elems * sizeof(T)));
  b.cap = elems;
  return DOAPP_OK;
}

template <typename T> void release(DevBuf<T> &b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
}

void invalidate_graph(doapp_ctx *ctx) {
  if (ctx->replan_exec) {
    cudaGraphExecDestroy(ctx->replan_exec);
    ctx->replan_exec = nullptr;
  }
  if (ctx->graph_alt.exec) {
    cudaGraphExecDestroy(ctx->graph_alt.exec);
    ctx->graph_alt.exec = nullptr;
  }
  if (ctx->graph_exec) {
    cudaGraphExecDestroy(ctx->graph_exec);
    ctx->graph_exec = nullptr;
  }
}

uint32_t f2u(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}

// Runs the exhaustive device check of dm::div_fast_unguarded against __fdiv_rn for divisor b
// over the floats with bit patterns [lo_bits, hi_bits]; *ok = no mismatch.
// The verdict depends only on (divisor, range, mode) -- the arithmetic is the same on every
// sm_100a device -- so it is cached per process: a caller that re-allocates its grid every
// frame pays the ~1 ms check once per resolution.
doapp_status verify_fast_div(doapp_ctx *ctx, float b, uint32_t lo_bits, uint32_t hi_bits, int mode,
                             bool *ok) {
  *ok = false;
  volatile float rcp = 1.0f / b;
  if (!ctx->fast_math_ok || !(b > 0x1p-60f && b < 0x1p+60f)) return DOAPP_OK;
  static std::mutex mu;
  static std::map<std::tuple<uint32_t, uint32_t, uint32_t, int>, bool> verdicts;
  const auto key = std::make_tuple(f2u(b), lo_bits, hi_bits, mode);
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = verdicts.find(key);
    if (it != verdicts.end()) {
      *ok = it->second;
      return DOAPP_OK;
    }
  }
  ST(ensure(ctx, ctx->vbad, 1));
  CU(cudaMemsetAsync(ctx->vbad.p, 0, sizeof(unsigned long long), ctx->stream));
  dm::FastDiv d{b, rcp, 0.0f, 0.0f};
  CU(launch_verify_div(d, lo_bits, hi_bits, mode, ctx->vbad.p, ctx->sm_count, ctx->stream));
  ctx->launches += 1;
  unsigned long long bad = 1;
  CU(cudaMemcpyAsync(&bad, ctx->vbad.p, sizeof(bad), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  *ok = bad == 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    verdicts[key] = *ok;
  }
  return DOAPP_OK;
}

CostParams base_cost_params(doapp_ctx *ctx) {
  const doapp_config &c = ctx->cfg;
  CostParams p{};
  p.traj = ctx->traj.p + (size_t)ctx->traj_s * ctx->traj_half;
  p.start = ctx->start.p;
  p.k = c.num_climbers;
  p.m = c.samples_per_climber;
  p.n = c.num_waypoints;
  p.npad = ctx->npad;
  p.n_noise = c.noise_waypoints;
  p.seed_lo = (uint32_t)c.seed;
  p.seed_hi = (uint32_t)(c.seed >> 32);
  p.sigma = c.sigma;
  p.sample_offset = c.sample_offset;
  p.m_global = ctx->m_global;
  p.dt = c.delta_t;
  p.smooth_scale = c.smooth_scale;
  p.coll_dist = c.collision_distance;
  p.limit_weight = c.limit_weight;
  for (int j = 0; j < kMaxJoints; ++j) {
    p.lim_min[j] = ctx->lim_min[j];
    p.lim_max[j] = ctx->lim_max[j];
  }
  p.use_limits = ctx->use_limits;
  p.collision_mode = c.collision_mode;
  p.select_mode = c.select_mode;
  p.rpy_s = ctx->rpy_s;
  p.rpy_c = ctx->rpy_c;
  p.fk_on = ctx->fk_on ? 1 : 0;
  if (ctx->fk_on) p.fk = ctx->fk_host;
  p.grid = ctx->grid;
  p.dt_div = ctx->dt_div;
  p.fast_ok = ctx->dt_fast_ok ? 1 : 0;
  p.pk_mode = ctx->pk_mode;
  return p;
}

float *traj_half_ptr(doapp_ctx *ctx, int half) {
  return ctx->traj.p + (size_t)half * ctx->traj_half;
}

void reset_plan_state(doapp_ctx *ctx) {
  ctx->traj_s = 0;
  ctx->cur_s = 0;
  ctx->key_w = -1;
}

// One cost launch of the planner sequence.  pend_mode != PEND_NONE: the launch first applies
// the keep-best decided by the key buffer the previous launch wrote (noise of `pend_iteration`
// / `pend_noise`), flipping the trajectory and cur_hi halves.  write_keys: the launch leaves
// its own per-climber winner in the next key buffer and re-arms the one after it.
doapp_status launch_pass(doapp_ctx *ctx, CostParams p, int pend_mode, uint32_t pend_iteration,
                         int pend_noise_mode, const float *pend_noise, bool write_keys) {
  const size_t k = ctx->cfg.num_climbers;
  p.traj = traj_half_ptr(ctx, ctx->traj_s);
  p.pend_mode = pend_mode;
  if (pend_mode != PEND_NONE) {
    if (ctx->key_w < 0) return fail(ctx, DOAPP_NOT_READY, "launch_pass: no pending keys");
    p.pend_keys = ctx->keys.p + (size_t)ctx->key_w * k;
    p.pend_iteration = pend_iteration;
    p.pend_noise_mode = pend_noise_mode;
    p.pend_noise = pend_noise;
    p.cur_hi_in = ctx->cur_hi.p + (size_t)ctx->cur_s * k;
    p.cur_hi_out = ctx->cur_hi.p + (size_t)(1 - ctx->cur_s) * k;
    p.found_better = ctx->found.p;
    p.traj_next = traj_half_ptr(ctx, 1 - ctx->traj_s);
  }
  int kw = -1;
  if (write_keys) {
    kw = (ctx->key_w + 1) % 3;
    p.keys = ctx->keys.p + (size_t)kw * k;
    p.keys_reset = ctx->keys.p + (size_t)((kw + 1) % 3) * k;
  }
  CU(launch_cost(p, ctx->cfg.waypoint_dim, ctx->sm_count, ctx->stream));
  if (pend_mode != PEND_NONE) {
    ctx->traj_s ^= 1;
    ctx->cur_s ^= 1;
  }
  if (write_keys) ctx->key_w = kw;
  return DOAPP_OK;
}

// cost pass over the k kept trajectories themselves (zero noise, one "sample"
// per climber): generate_trajectories.cu:180-191.
CostParams climber_cost_params(doapp_ctx *ctx) {
  CostParams p = base_cost_params(ctx);
  p.m = 1;
  p.noise_mode = NOISE_ZERO;
  p.noise = nullptr;
  p.n_noise = ctx->cfg.noise_waypoints; // layout of a PENDING winner's noise; own noise is zero
  p.sample_offset = 0;
  p.m_global = 1;
  return p;
}

doapp_status upload_traj(doapp_ctx *ctx, const float *host_aos, bool broadcast_first) {
  const doapp_config &c = ctx->cfg;
  size_t rows = broadcast_first ? 1 : c.num_climbers;
  size_t elems = rows * c.num_waypoints * c.waypoint_dim;
  ST(ensure(ctx, ctx->aos_tmp, (size_t)c.num_climbers * c.num_waypoints * c.waypoint_dim));
  CU(cudaMemcpyAsync(ctx->aos_tmp.p, host_aos, elems * sizeof(float), cudaMemcpyHostToDevice,
                     ctx->stream));
  reset_plan_state(ctx); // the uploaded trajectories live in half 0
  CU(launch_aos_to_soa(ctx->aos_tmp.p, ctx->traj.p, c.num_climbers, c.num_waypoints, ctx->npad,
                       c.waypoint_dim, broadcast_first ? 1 : 0, ctx->stream));
  ctx->launches += 1;
  return DOAPP_OK;
}

doapp_status upload_start(doapp_ctx *ctx, const float *start) {
  CU(cudaMemcpyAsync(ctx->start.p, start, ctx->cfg.waypoint_dim * sizeof(float),
                     cudaMemcpyHostToDevice, ctx->stream));
  return DOAPP_OK;
}

// The whole keep-best loop in one cooperative launch (plan_persistent_kernel) when the problem
// fits one co-resident wave: single rank, keep-best rule, one noise vector per sample.
doapp_status persistent_eligible(doapp_ctx *ctx, PersistParams *q, bool *ok) {
  *ok = false;
  const doapp_config &c = ctx->cfg;
  if (ctx->persist_mode == 0 || c.update_mode != DOAPP_UPDATE_KEEP_BEST || c.noise_waypoints != 1)
    return DOAPP_OK;
  // First generation: ONE rank only.  Its in-kernel mailbox exchange polls without a bound
  // (ADVICE r01); several ranks take plan_pipeline_kernel (bounded waits, error word) or the
  // launch-per-iteration planner with xchg_send_kernel / NCCL instead.
  if (ctx->comm && ctx->nranks > 1) return DOAPP_OK;
  CostParams p = base_cost_params(ctx);
  CU(persistent_plan_config(p, c.waypoint_dim, ctx->sm_count, q, ok));
  return DOAPP_OK;
}

// Second-generation persistent planner (plan_pipeline_kernel); same rules, same rank-uniform
// verdict.  Preferred over the first generation whenever it is eligible.
doapp_status pipeline_eligible(doapp_ctx *ctx, Persist2Params *q, bool *ok) {
  *ok = false;
  const doapp_config &c = ctx->cfg;
  if (ctx->persist_mode == 0 || ctx->persist_version != 2 ||
      c.update_mode != DOAPP_UPDATE_KEEP_BEST || c.noise_waypoints != 1)
    return DOAPP_OK;
  if (ctx->comm && !(ctx->peer_ready && ctx->peer_nranks == ctx->nranks)) return DOAPP_OK;
  CostParams p = base_cost_params(ctx);
  if (ctx->comm && ctx->nranks > 1) {
    CostParams pm = p;
    pm.m = (uint32_t)((ctx->m_global + ctx->nranks - 1) / ctx->nranks);
    pm.sample_offset = ctx->m_global - pm.m; // the last shard: largest offset
    Persist2Params qm{};
    CU(pipeline_plan_config(pm, c.waypoint_dim, ctx->sm_count, &qm, ok));
    if (!*ok) return DOAPP_OK;
  }
  CU(pipeline_plan_config(p, c.waypoint_dim, ctx->sm_count, q, ok));
  return DOAPP_OK;
}

doapp_status enqueue_final(doapp_ctx *ctx) {
  const doapp_config &c = ctx->cfg;
  CostParams f0 = climber_cost_params(ctx);
  f0.scores = ctx->final_scores.p;
  f0.collisions = ctx->final_coll.p;
  ST(launch_pass(ctx, f0, PEND_NONE, 0, NOISE_ZERO, nullptr, false));
  FinalParams f{};
  f.traj = traj_half_ptr(ctx, ctx->traj_s);
  f.start = ctx->start.p;
  f.scores = ctx->final_scores.p;
  f.found_better = ctx->found.p;
  f.k = c.num_climbers;
  f.n = c.num_waypoints;
  f.npad = ctx->npad;
  f.d = c.waypoint_dim;
  f.dt = c.delta_t;
  f.out = ctx->out.p;
  CU(launch_final(f, ctx->stream));
  return DOAPP_OK;
}

doapp_status enqueue_plan_pipeline(doapp_ctx *ctx, Persist2Params q, uint32_t iters, int noise_mode,
                                   uint64_t *kernels) {
  const doapp_config &c = ctx->cfg;
  const size_t k = c.num_climbers;
  reset_plan_state(ctx);
  CU(cudaMemsetAsync(ctx->keys.p, 0xFF, sizeof(unsigned long long) * 3 * k, ctx->stream));
  // seed pass: score the k unperturbed trajectories (gt.cu:180-191); its keys seed cur_hi
  ST(launch_pass(ctx, climber_cost_params(ctx), PEND_NONE, 0, NOISE_ZERO, nullptr, true));
  const size_t nent = (size_t)kP2Ring * k * pipeline_grid_x(q);
  ST(ensure(ctx, ctx->p2_entries, nent));
  CU(cudaMemsetAsync(ctx->p2_entries.p, 0xFF, sizeof(unsigned long long) * nent, ctx->stream));
  CostParams &p = q.cost;
  p.traj = traj_half_ptr(ctx, ctx->traj_s);
  p.traj_next = traj_half_ptr(ctx, 1 - ctx->traj_s);
  p.pend_keys = ctx->keys.p + (size_t)ctx->key_w * k;
  p.cur_hi_out = ctx->cur_hi.p + (size_t)(1 - ctx->cur_s) * k;
  p.found_better = ctx->found.p;
  p.noise_mode = noise_mode;
  p.noise_m = ctx->m_global;
  p.noise_t0 = 0;
  q.iters = iters;
  q.entries = ctx->p2_entries.p;
  q.noise_base = noise_mode == NOISE_EXTERNAL ? ctx->noise.p : nullptr;
  q.noise_iter_stride = (uint64_t)c.num_climbers * ctx->m_global * c.waypoint_dim;
  q.nranks = 1;
  q.rank = 0;
  q.error = ctx->p2_error.p;
  q.stats = ctx->p2_stats.p;
  q.timeout_ns = 20ull * 1000000000ull; // a peer that never arrives: leave after 20 s
  {
    static const char *pv = getenv("DOAPP_P2_POLL_NS"); // tuning
    q.poll_ns = pv ? (uint32_t)atoi(pv) : 100u;
  }
  ctx->last_plan_mode = 1;
  const bool multi = ctx->comm && ctx->peer_ready;
  if (multi) {
    ctx->last_plan_mode = 3;
    const size_t set = (size_t)kP2Ring * k * 8;
    q.nranks = (uint32_t)ctx->peer_nranks;
    q.rank = (uint32_t)ctx->peer_rank;
    for (int r = 0; r < ctx->peer_nranks; ++r) q.mailbox[r] = ctx->peer_mbox[r] + (ctx->peer_epoch & 1) * set;
    q.mailbox_next = ctx->mailbox.p + ((ctx->peer_epoch + 1) & 1) * set;
    ctx->peer_epoch++;
  }
  q.depth = ctx->persist_depth ? (uint32_t)ctx->persist_depth : (multi ? 10u : 8u);
  if (q.depth > (uint32_t)kP2MaxDepth) q.depth = kP2MaxDepth;
  if (ctx->profiling) CU(cudaEventRecord(ctx->prof_events[0], ctx->stream));
  CU(launch_pipeline_plan(q, c.waypoint_dim, ctx->sm_count, ctx->stream));
  if (ctx->profiling) CU(cudaEventRecord(ctx->prof_events[1], ctx->stream));
  ctx->prof_launches = 1;
  ctx->p2_ran = true;
  ctx->traj_s ^= 1;
  ctx->cur_s ^= 1;
  ST(enqueue_final(ctx));
  if (kernels) *kernels = 4;
  return DOAPP_OK;
}

doapp_status enqueue_plan_persistent(doapp_ctx *ctx, PersistParams q, uint32_t iters, int noise_mode,
                                     uint64_t *kernels) {
  const doapp_config &c = ctx->cfg;
  const size_t k = c.num_climbers;
  reset_plan_state(ctx);
  CU(cudaMemsetAsync(ctx->keys.p, 0xFF, sizeof(unsigned long long) * 3 * k, ctx->stream));
  // seed pass: score the k unperturbed trajectories (gt.cu:180-191); its keys seed cur_hi
  ST(launch_pass(ctx, climber_cost_params(ctx), PEND_NONE, 0, NOISE_ZERO, nullptr, true));
  ST(ensure(ctx, ctx->pkeys, 3 * k * 2));
  CU(cudaMemsetAsync(ctx->pkeys.p, 0xFF, sizeof(unsigned long long) * 3 * k * 2, ctx->stream));
  CostParams &p = q.cost;
  p.traj = traj_half_ptr(ctx, ctx->traj_s);
  p.traj_next = traj_half_ptr(ctx, 1 - ctx->traj_s);
  p.pend_keys = ctx->keys.p + (size_t)ctx->key_w * k;
  p.cur_hi_out = ctx->cur_hi.p + (size_t)(1 - ctx->cur_s) * k;
  p.found_better = ctx->found.p;
  p.noise_mode = noise_mode;
  p.noise_m = ctx->m_global;
  p.noise_t0 = 0;
  q.iters = iters;
  q.keys3 = ctx->pkeys.p;
  q.noise_base = noise_mode == NOISE_EXTERNAL ? ctx->noise.p : nullptr;
  q.noise_iter_stride = (uint64_t)c.num_climbers * ctx->m_global * c.waypoint_dim;
  q.nranks = 1;
  q.rank = 0;
  ctx->last_plan_mode = 1;
  if (ctx->comm && ctx->peer_ready) {
    ctx->last_plan_mode = 3;
    // mailbox set of this launch (sets alternate per launch; the kernel re-arms the other one)
    const size_t set = 3 * k * 8;
    q.nranks = (uint32_t)ctx->peer_nranks;
    q.rank = (uint32_t)ctx->peer_rank;
    for (int r = 0; r < ctx->peer_nranks; ++r) q.mailbox[r] = ctx->peer_mbox[r] + (ctx->peer_epoch & 1) * set;
    q.mailbox_next = ctx->mailbox.p + ((ctx->peer_epoch + 1) & 1) * set;
    q.result = ctx->presult.p;
    CU(cudaMemsetAsync(ctx->presult.p, 0xFF, sizeof(unsigned long long) * 3 * k, ctx->stream));
    ctx->peer_epoch++;
  }
  if (ctx->profiling) CU(cudaEventRecord(ctx->prof_events[0], ctx->stream));
  CU(launch_persistent_plan(q, c.waypoint_dim, ctx->stream));
  if (ctx->profiling) CU(cudaEventRecord(ctx->prof_events[1], ctx->stream));
  ctx->prof_launches = 1;
  ctx->traj_s ^= 1;
  ctx->cur_s ^= 1;
  {
    CostParams f0 = climber_cost_params(ctx);
    f0.scores = ctx->final_scores.p;
    f0.collisions = ctx->final_coll.p;
    ST(launch_pass(ctx, f0, PEND_NONE, 0, NOISE_ZERO, nullptr, false));
    FinalParams f{};
    f.traj = traj_half_ptr(ctx, ctx->traj_s);
    f.start = ctx->start.p;
    f.scores = ctx->final_scores.p;
    f.found_better = ctx->found.p;
    f.k = c.num_climbers;
    f.n = c.num_waypoints;
    f.npad = ctx->npad;
    f.d = c.waypoint_dim;
    f.dt = c.delta_t;
    f.out = ctx->out.p;
    CU(launch_final(f, ctx->stream));
  }
  if (kernels) *kernels = 4;
  return DOAPP_OK;
}

// Enqueue the planner loop (no host synchronisation inside).  One launch per stage:
//   seed pass   scores the k unperturbed trajectories                      (gt.cu:180-191)
//   iteration i costs the samples; while staging it first applies the keep-best of
//               iteration i-1 (or seeds cur_hi from the seed pass)          (gt.cu:197-241)
//   [ranks > 1: one MIN all-reduce of the k keys after every iteration      ]
//   final pass  applies the last keep-best and scores the kept trajectories in float
//   selection   host-style argmin + velocities                              (ps.cu:113-166)
doapp_status enqueue_plan(doapp_ctx *ctx, uint32_t iters, int noise_mode, uint64_t *kernels) {
  const doapp_config &c = ctx->cfg;
  const size_t k = c.num_climbers;
  uint64_t nk = 0;
  NcclApi *nccl = ctx->comm ? nccl_api(nullptr) : nullptr;
  ctx->p2_ran = false;
  {
    Persist2Params q2{};
    bool pipeline = false;
    ST(pipeline_eligible(ctx, &q2, &pipeline));
    if (pipeline) return enqueue_plan_pipeline(ctx, q2, iters, noise_mode, kernels);
    PersistParams q{};
    bool persistent = false;
    ST(persistent_eligible(ctx, &q, &persistent));
    if (persistent) return enqueue_plan_persistent(ctx, q, iters, noise_mode, kernels);
  }
  ctx->prof_launches = iters;
  ctx->last_plan_mode = 0;
  reset_plan_state(ctx);
  CU(cudaMemsetAsync(ctx->keys.p, 0xFF, sizeof(unsigned long long) * 3 * k, ctx->stream));
  ST(launch_pass(ctx, climber_cost_params(ctx), PEND_NONE, 0, NOISE_ZERO, nullptr, true));
  nk += 1;
  // Several ranks with the peer mailboxes mapped: the per-iteration MIN all-reduce becomes one
  // tiny sending kernel after every cost launch (8-byte stores into every rank's mailbox over
  // NVLink) and a poll of the own mailbox in the next cost launch's prologue -- no collective
  // launch, no NCCL kernel, in the loop.  Mailbox sets alternate per plan (xr_parity).
  const bool xr = ctx->xr_on && nccl != nullptr;
  const size_t xr_slot = k * 8, xr_set = 3 * xr_slot;
  unsigned long long *xr_own = xr ? ctx->mailbox.p + (size_t)ctx->xr_parity * xr_set : nullptr;
  auto xr_recv_from = [&](CostParams &p, uint32_t it_of_keys) {
    p.xr_nranks = (uint32_t)ctx->peer_nranks;
    p.xr_recv = xr_own + (size_t)(it_of_keys % 3) * xr_slot;
    p.xr_error = ctx->xr_error.p;
    p.xr_timeout_ns = 20ull * 1000000000ull;
  };
  if (xr) ctx->last_plan_mode = 2;
  const size_t noise_iter_stride =
      (size_t)c.num_climbers * ctx->m_global * c.noise_waypoints * c.waypoint_dim;
  auto noise_of = [&](uint32_t it) -> const float * {
    return noise_mode == NOISE_EXTERNAL ? ctx->noise.p + (size_t)it * noise_iter_stride : nullptr;
  };
  const bool weighted = c.update_mode == DOAPP_UPDATE_WEIGHTED;
  const int pend_it = weighted ? PEND_WEIGHTED : PEND_KEEP_BEST;
  if (weighted)
    CU(cudaMemsetAsync(ctx->usum.p, 0, sizeof(long long) * 2 * k * 8, ctx->stream));
  for (uint32_t it = 0; it < iters; ++it) {
    CostParams p = base_cost_params(ctx);
    p.noise_mode = noise_mode;
    p.iteration = it;
    p.noise = noise_of(it);
    p.noise_m = ctx->m_global;
    p.noise_t0 = 0;
    if (weighted) {
      // the update vector needs every sample's score and the block's int(ceilf(.)) minimum
      p.scores = ctx->scores.p;
      p.select_mode = DOAPP_SELECT_CEIL_INT;
      p.update_gain = c.update_gain;
      p.pend_usum = ctx->usum.p + (size_t)((it + 1) & 1) * k * 8; // sums of iteration it-1
      p.usum_reset = ctx->usum.p + (size_t)(it & 1) * k * 8;      // zeroed for this iteration
    }
    if (xr && it > 0) xr_recv_from(p, it - 1);
    if (ctx->profiling) CU(cudaEventRecord(ctx->prof_events[2 * it], ctx->stream));
    ST(launch_pass(ctx, p, it == 0 ? PEND_SEED : pend_it, it ? it - 1 : 0, noise_mode,
                   it ? noise_of(it - 1) : nullptr, true));
    if (ctx->profiling) CU(cudaEventRecord(ctx->prof_events[2 * it + 1], ctx->stream));
    nk += 1;
    if (xr) {
      XchgSendParams x{};
      x.keys = ctx->keys.p + (size_t)ctx->key_w * k;
      for (int r = 0; r < ctx->peer_nranks; ++r)
        x.peer_slot[r] = ctx->peer_mbox[r] + (size_t)ctx->xr_parity * xr_set + (size_t)(it % 3) * xr_slot;
      x.rearm_slot = xr_own + (size_t)((it + 1) % 3) * xr_slot;
      x.rearm_set = it == 0 ? ctx->mailbox.p + (size_t)(1 - ctx->xr_parity) * xr_set : nullptr;
      x.k = (uint32_t)k;
      x.nranks = (uint32_t)ctx->peer_nranks;
      x.rank = (uint32_t)ctx->peer_rank;
      CU(launch_xchg_send(x, ctx->stream));
      nk += 1;
    } else if (nccl) {
      unsigned long long *kb = ctx->keys.p + (size_t)ctx->key_w * k;
      ncclResult_t r = nccl->AllReduce(kb, kb, k, ncclUint64, ncclMin, ctx->comm, ctx->stream);
      if (r != ncclSuccess)
        return fail(ctx, DOAPP_NCCL_ERROR, "ncclAllReduce: %s",
                    nccl->GetErrorString ? nccl->GetErrorString(r) : "error");
    }
    if (weighted) {
      // compute_update_vector (gt.cu:117-133) over this rank's samples, integer sums
      UpdateParams u{};
      u.scores = ctx->scores.p;
      u.keys = ctx->keys.p + (size_t)ctx->key_w * k;
      u.noise = noise_of(it);
      u.noise_m = ctx->m_global;
      u.noise_t0 = 0;
      u.noise_mode = noise_mode;
      u.k = c.num_climbers;
      u.m = c.samples_per_climber;
      u.d = c.waypoint_dim;
      u.iteration = it;
      u.seed_lo = (uint32_t)c.seed;
      u.seed_hi = (uint32_t)(c.seed >> 32);
      u.sigma = c.sigma;
      u.sample_offset = c.sample_offset;
      u.usum = ctx->usum.p + (size_t)(it & 1) * k * 8;
      CU(launch_update_vector(u, ctx->sm_count, ctx->stream));
      nk += 1;
      if (nccl) { // exact: integer sums commute
        ncclResult_t r =
            nccl->AllReduce(u.usum, u.usum, k * 8, ncclInt64, ncclSum, ctx->comm, ctx->stream);
        if (r != ncclSuccess)
          return fail(ctx, DOAPP_NCCL_ERROR, "ncclAllReduce(sum): %s",
                      nccl->GetErrorString ? nccl->GetErrorString(r) : "error");
      }
    }
  }
  {
    CostParams p = climber_cost_params(ctx);
    p.scores = ctx->final_scores.p;
    p.collisions = ctx->final_coll.p;
    if (weighted) {
      p.update_gain = c.update_gain;
      p.pend_usum = ctx->usum.p + (size_t)((iters + 1) & 1) * k * 8;
    }
    // the winner's noise of the last iteration is still needed to apply its keep-best
    p.noise_m = ctx->m_global;
    p.noise_t0 = 0;
    if (xr && iters > 0) xr_recv_from(p, iters - 1);
    ST(launch_pass(ctx, p, iters == 0 ? PEND_SEED : pend_it, iters ? iters - 1 : 0,
                   noise_mode, iters ? noise_of(iters - 1) : nullptr, false));
    FinalParams f{};
    f.traj = traj_half_ptr(ctx, ctx->traj_s);
    f.start = ctx->start.p;
    f.scores = ctx->final_scores.p;
    f.found_better = ctx->found.p;
    f.k = c.num_climbers;
    f.n = c.num_waypoints;
    f.npad = ctx->npad;
    f.d = c.waypoint_dim;
    f.dt = c.delta_t;
    f.out = ctx->out.p;
    CU(launch_final(f, ctx->stream));
    nk += 2;
  }
  if (kernels) *kernels = nk;
  return DOAPP_OK;
}

// restore every climber to the staged initial trajectory (device copy kept in aos_tmp)
doapp_status reset_climbers(doapp_ctx *ctx) {
  const doapp_config &c = ctx->cfg;
  reset_plan_state(ctx);
  CU(launch_aos_to_soa(ctx->aos_tmp.p, ctx->traj.p, c.num_climbers, c.num_waypoints, ctx->npad,
                       c.waypoint_dim, 1, ctx->stream));
  ctx->launches += 1;
  return DOAPP_OK;
}

doapp_status run_plan(doapp_ctx *ctx, uint32_t iters, int noise_mode) {
  PersistParams pq{};
  bool persistent = false;
  ST(persistent_eligible(ctx, &pq, &persistent));
  bool pipeline = false;
  {
    Persist2Params pq2{};
    ST(pipeline_eligible(ctx, &pq2, &pipeline));
    if (pipeline) // allocate outside any capture
      ST(ensure(ctx, ctx->p2_entries, (size_t)kP2Ring * ctx->cfg.num_climbers * pipeline_grid_x(pq2)));
  }
  // The persistent planners are a handful of launches, but their launch gaps (memsets, seed
  // pass, planner, final pass, selection) still cost ~0.2 us per iteration of a 100-iteration
  // plan: the single-rank pipeline planner is replayed as a graph too.  Several ranks: the
  // mailbox set alternates per launch, so the sequence is enqueued directly.
  const bool graph_ok = ctx->use_graph && !ctx->profiling &&
                        (pipeline ? (!ctx->comm && ctx->persist_graph_ok) : !persistent);
  // launch-per-iteration path on several ranks: keys through the peer mailboxes when they are
  // mapped (keep-best rule; the weighted rule also needs its exact integer SUM: NCCL)
  ctx->xr_on = !pipeline && !persistent && ctx->comm && ctx->peer_ready &&
               ctx->peer_nranks == ctx->nranks && ctx->xr_mode != 0 &&
               ctx->cfg.update_mode == DOAPP_UPDATE_KEEP_BEST;
  if (ctx->xr_on) {
    ctx->xr_parity = (int)(ctx->peer_epoch & 1);
    ctx->peer_epoch++;
    if (ctx->graph_parity != ctx->xr_parity) { // the other mailbox set: the other cached graph
      doapp_ctx::GraphAlt cur;
      cur.exec = ctx->graph_exec; cur.iters = ctx->graph_iters; cur.noise_mode = ctx->graph_noise_mode;
      cur.noise_ptr = ctx->graph_noise_ptr; cur.grid_generation = ctx->graph_grid_generation;
      cur.kernels = ctx->graph_kernels; cur.end_traj_s = ctx->graph_end_traj_s;
      cur.end_cur_s = ctx->graph_end_cur_s; cur.end_key_w = ctx->graph_end_key_w;
      const doapp_ctx::GraphAlt alt = ctx->graph_alt;
      ctx->graph_exec = alt.exec; ctx->graph_iters = alt.iters; ctx->graph_noise_mode = alt.noise_mode;
      ctx->graph_noise_ptr = alt.noise_ptr; ctx->graph_grid_generation = alt.grid_generation;
      ctx->graph_kernels = alt.kernels; ctx->graph_end_traj_s = alt.end_traj_s;
      ctx->graph_end_cur_s = alt.end_cur_s; ctx->graph_end_key_w = alt.end_key_w;
      ctx->graph_alt = cur;
      ctx->graph_parity = ctx->xr_parity;
    }
  }
  if (!graph_ok) {
    uint64_t nk = 0;
    ST(enqueue_plan(ctx, iters, noise_mode, &nk));
    ctx->launches += nk;
    return DOAPP_OK;
  }
  const float *nptr = noise_mode == NOISE_EXTERNAL ? ctx->noise.p : nullptr;
  bool hit = ctx->graph_exec && ctx->graph_iters == iters && ctx->graph_noise_mode == noise_mode &&
             ctx->graph_noise_ptr == nptr && ctx->graph_grid_generation == ctx->grid_generation;
  if (!hit) {
    if (ctx->graph_exec) { // this parity's graph only; the other mailbox set's graph stays cached
      cudaGraphExecDestroy(ctx->graph_exec);
      ctx->graph_exec = nullptr;
    }
    if (ctx->replan_exec) {
      cudaGraphExecDestroy(ctx->replan_exec);
      ctx->replan_exec = nullptr;
    }
    cudaGraph_t graph = nullptr;
    CU(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    uint64_t nk = 0;
    doapp_status st = enqueue_plan(ctx, iters, noise_mode, &nk);
    cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
    if (ce == cudaSuccess && st == DOAPP_OK) {
      ce = cudaGraphInstantiate(&ctx->graph_exec, graph, 0);
      if (ce != cudaSuccess) ctx->graph_exec = nullptr;
    }
    if (graph) cudaGraphDestroy(graph);
    if (pipeline && (st != DOAPP_OK || ce != cudaSuccess)) {
      // a cooperative launch that cannot be captured / instantiated here: enqueue directly
      cudaGetLastError();
      ctx->persist_graph_ok = false;
      uint64_t nk2 = 0;
      ST(enqueue_plan(ctx, iters, noise_mode, &nk2));
      ctx->launches += nk2;
      return DOAPP_OK;
    }
    if (st != DOAPP_OK) return st;
    if (ce != cudaSuccess)
      return fail(ctx, DOAPP_CUDA_ERROR, "graph capture / instantiate: %s", cudaGetErrorString(ce));
    ctx->graph_iters = iters;
    ctx->graph_noise_mode = noise_mode;
    ctx->graph_noise_ptr = nptr;
    ctx->graph_grid_generation = ctx->grid_generation;
    ctx->graph_kernels = nk;
    ctx->graph_end_traj_s = ctx->traj_s;
    ctx->graph_end_cur_s = ctx->cur_s;
    ctx->graph_end_key_w = ctx->key_w;
  }
  CU(cudaGraphLaunch(ctx->graph_exec, ctx->stream));
  ctx->traj_s = ctx->graph_end_traj_s; // where the replayed sequence leaves the ping-pong state
  ctx->cur_s = ctx->graph_end_cur_s;
  ctx->key_w = ctx->graph_end_key_w;
  ctx->launches += ctx->graph_kernels;
  return DOAPP_OK;
}

} // namespace

// ======================================================================= ABI

extern "C" {

void doapp_config_init(doapp_config *cfg) {
  if (!cfg) return;
  memset(cfg, 0, sizeof(*cfg));
  cfg->struct_size = sizeof(doapp_config);
  cfg->device = 0;
  cfg->num_climbers = 1;
  cfg->samples_per_climber = 1024;
  cfg->num_waypoints = 64;
  cfg->waypoint_dim = 5;
  cfg->noise_waypoints = 1;
  cfg->delta_t = 0.1f;
  cfg->smooth_scale = 0.25f;
  cfg->collision_distance = 0.05f;
  cfg->collision_mode = DOAPP_COLLISION_INTENDED;
  cfg->select_mode = DOAPP_SELECT_CEIL_INT;
  cfg->update_mode = DOAPP_UPDATE_KEEP_BEST;
  cfg->update_gain = 10000.0f;
  cfg->limit_weight = 0.0f;
  cfg->sigma = 1.0f;
  cfg->seed = 1234;
  cfg->sample_offset = 0;
  cfg->global_samples_per_climber = 0;
}

uint32_t doapp_version(void) { return DOAPP_ABI_VERSION; }

const char *doapp_status_str(doapp_status s) {
  switch (s) {
  case DOAPP_OK: return "OK";
  case DOAPP_INVALID_ARG: return "INVALID_ARG";
  case DOAPP_CUDA_ERROR: return "CUDA_ERROR";
  case DOAPP_NCCL_ERROR: return "NCCL_ERROR";
  case DOAPP_OOM: return "OOM";
  case DOAPP_NOT_READY: return "NOT_READY";
  }
  return "UNKNOWN";
}

const char *doapp_last_error(const doapp_ctx *ctx) {
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

doapp_status doapp_create(const doapp_config *cfg, doapp_ctx **out) {
  doapp_ctx *ctx = nullptr; // for the CU/fail macros: create errors go to the thread-local slot
  if (!cfg || !out) return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: null argument");
  *out = nullptr;
  if (cfg->struct_size != sizeof(doapp_config))
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: struct_size %u != %zu",
                cfg->struct_size, sizeof(doapp_config));
  if (cfg->num_climbers == 0 || cfg->samples_per_climber == 0)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: k and m must be > 0");
  if (cfg->num_climbers > 65535)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: num_climbers must be <= 65535");
  if (cfg->num_waypoints == 0 || cfg->num_waypoints > 1024)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: num_waypoints must be in 1..1024");
  if (cfg->waypoint_dim != 4 && cfg->waypoint_dim != 5)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: waypoint_dim must be 4 or 5");
  if (cfg->noise_waypoints != 1 && cfg->noise_waypoints != cfg->num_waypoints)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: noise_waypoints must be 1 or n");
  if (!(cfg->delta_t > 0.0f))
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: delta_t must be > 0");
  if (cfg->update_mode != DOAPP_UPDATE_KEEP_BEST && cfg->update_mode != DOAPP_UPDATE_WEIGHTED)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: unknown update_mode %d",
                cfg->update_mode);
  uint64_t mg = cfg->global_samples_per_climber ? cfg->global_samples_per_climber
                                                : cfg->samples_per_climber;
  // weighted update: one d-vector of noise per sample (generate_trajectories.cu:117-142), and the
  // 64-bit fixed-point sums (terms < 2^12 scaled by 2^28) must not overflow
  if (cfg->update_mode == DOAPP_UPDATE_WEIGHTED && (cfg->noise_waypoints != 1 || mg > (1ull << 22)))
    return fail(nullptr, DOAPP_INVALID_ARG,
                "doapp_create: DOAPP_UPDATE_WEIGHTED needs noise_waypoints == 1 and m_global <= 2^22");
  if (cfg->sample_offset + cfg->samples_per_climber > mg || mg > 0xFFFFFFFFull)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: shard [%llu,+%u) outside m_global %llu",
                (unsigned long long)cfg->sample_offset, cfg->samples_per_climber,
                (unsigned long long)mg);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, DOAPP_CUDA_ERROR,
                "doapp_create: no CUDA device (%s); this library has no CPU fallback",
                cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= ndev)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_create: device %d of %d", cfg->device, ndev);
  CU(cudaSetDevice(cfg->device));
  ctx = new (std::nothrow) doapp_ctx();
  if (!ctx) return fail(nullptr, DOAPP_OOM, "doapp_create: host allocation failed");
  ctx->cfg = *cfg;
  ctx->m_global = mg;
  ctx->npad = (cfg->num_waypoints + 31u) & ~31u;
  ctx->S = (uint64_t)cfg->num_climbers * cfg->samples_per_climber;
  const char *ng = getenv("DOAPP_NO_GRAPH");
  ctx->use_graph = !(ng && ng[0] == '1');
  const char *psv = getenv("DOAPP_PERSIST");
  ctx->persist_mode = (psv && (psv[0] == '0' || psv[0] == '1')) ? psv[0] - '0' : -1;
  const char *pvv = getenv("DOAPP_PERSIST_V");
  ctx->persist_version = (pvv && pvv[0] == '1') ? 1 : 2;
  const char *pdv = getenv("DOAPP_PERSIST_DEPTH");
  ctx->persist_depth = pdv ? atoi(pdv) : 0;
  if (ctx->persist_depth < 0 || ctx->persist_depth > kP2MaxDepth) ctx->persist_depth = 0;
  const char *xcv = getenv("DOAPP_XCHG");
  ctx->xr_mode = (xcv && xcv[0] == '0') ? 0 : -1;
  const char *pkv = getenv("DOAPP_PK");
  ctx->pk_mode = (pkv && (pkv[0] == '0' || pkv[0] == '1')) ? pkv[0] - '0' : -1;
  const char *xd = getenv("DOAPP_EXACT_DIV");
  ctx->fast_math_ok = !(xd && xd[0] == '1');
  host_sincos(-1.5707963267948966f, &ctx->rpy_s, &ctx->rpy_c);
  doapp_status st = [&]() -> doapp_status {
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device));
    ctx->sm_count = prop.multiProcessorCount;
    CU(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&ctx->ev0));
    CU(cudaEventCreate(&ctx->ev1));
    const size_t k = cfg->num_climbers, n = cfg->num_waypoints, d = cfg->waypoint_dim;
    ctx->traj_half = k * d * ctx->npad;
    ST(ensure(ctx, ctx->traj, 2 * ctx->traj_half));
    ST(ensure(ctx, ctx->aos_tmp, k * n * d));
    ST(ensure(ctx, ctx->start, 8));
    ST(ensure(ctx, ctx->scores, ctx->S));
    ST(ensure(ctx, ctx->coll, ctx->S));
    ST(ensure(ctx, ctx->final_scores, k));
    ST(ensure(ctx, ctx->final_coll, k));
    ST(ensure(ctx, ctx->keys, 3 * k));
    ST(ensure(ctx, ctx->keys_float, k));
    ST(ensure(ctx, ctx->cur_hi, 2 * k));
    ST(ensure(ctx, ctx->usum, 2 * k * 8));
    ST(ensure(ctx, ctx->found, k));
    ST(ensure(ctx, ctx->out, 4 + 2 * n * d));
    ST(ensure(ctx, ctx->xr_error, 1));
    CU(cudaMemsetAsync(ctx->xr_error.p, 0, sizeof(unsigned int), ctx->stream));
    ST(ensure(ctx, ctx->p2_error, 1));
    ST(ensure(ctx, ctx->p2_stats, 8 + 256 * 8));
    CU(cudaMemsetAsync(ctx->p2_error.p, 0, sizeof(unsigned int), ctx->stream));
    CU(cudaMemsetAsync(ctx->p2_stats.p, 0, (8 + 256 * 8) * sizeof(unsigned long long), ctx->stream));
    CU(cudaMemsetAsync(ctx->traj.p, 0, sizeof(float) * 2 * ctx->traj_half, ctx->stream));
    CU(cudaMemsetAsync(ctx->start.p, 0, sizeof(float) * 8, ctx->stream));
    CU(cudaMemsetAsync(ctx->found.p, 0, k, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    // velocity / acceleration divides: reciprocal-multiply path, proven against the IEEE
    // divide for EVERY dividend in [2^-60, 2^60) for this delta_t, else disabled
    bool ok = false;
    if (cfg->delta_t >= dm::kFastDtLo && cfg->delta_t <= dm::kFastDtHi)
      ST(verify_fast_div(ctx, cfg->delta_t, f2u(dm::kFastDivLo), f2u(dm::kFastDivHi) - 1, 0, &ok));
    ctx->dt_fast_ok = ok;
    volatile float rcp = 1.0f / cfg->delta_t;
    ctx->dt_div = dm::FastDiv{cfg->delta_t, rcp, ok ? dm::kFastDivLo : INFINITY,
                              ok ? dm::kFastDivHi : INFINITY};
    return DOAPP_OK;
  }();
  if (st != DOAPP_OK) {
    g_create_error = ctx->err;
    doapp_destroy(ctx);
    return st;
  }
  ctx->h_out.resize(4 + 2 * (size_t)cfg->num_waypoints * cfg->waypoint_dim);
  *out = ctx;
  return DOAPP_OK;
}

void doapp_destroy(doapp_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->cfg.device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  invalidate_graph(ctx);
  if (ctx->comm) {
    NcclApi *api = nccl_api(nullptr);
    if (api && api->CommDestroy) api->CommDestroy(ctx->comm);
  }
  release(ctx->traj); release(ctx->aos_tmp); release(ctx->start); release(ctx->noise);
  release(ctx->scores); release(ctx->final_scores); release(ctx->out); release(ctx->cloud);
  release(ctx->pts); release(ctx->coll); release(ctx->final_coll); release(ctx->keys);
  release(ctx->keys_float); release(ctx->cur_hi); release(ctx->found); release(ctx->grid_cells);
  release(ctx->lookup_coll); release(ctx->lookup_idx); release(ctx->dbg_a); release(ctx->dbg_b);
  release(ctx->dbg_c); release(ctx->vbad); release(ctx->usum); release(ctx->pkeys);
  release(ctx->p2_entries); release(ctx->p2_stats); release(ctx->p2_error); release(ctx->xr_error);
  for (int r = 0; r < 8; ++r)
    if (ctx->peer_ready && r != ctx->peer_rank && ctx->peer_mbox[r]) cudaIpcCloseMemHandle(ctx->peer_mbox[r]);
  release(ctx->mailbox); release(ctx->presult); release(ctx->depth); release(ctx->fk_dev);
  for (cudaEvent_t e : ctx->prof_events) cudaEventDestroy(e);
  for (cudaEvent_t e : ctx->cloud_events) cudaEventDestroy(e);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

// ------------------------------------------------------------------ world model

doapp_status doapp_grid_alloc(doapp_ctx *ctx, int32_t kind, int32_t frame, uint32_t nx,
                              uint32_t ny, uint32_t nz, double resolution) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (kind != DOAPP_GRID_U8 && kind != DOAPP_GRID_F32)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_alloc: kind %d", kind);
  if (frame != DOAPP_FRAME_CDG && frame != DOAPP_FRAME_OG)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_alloc: frame %d", frame);
  if (nx == 0 || ny == 0 || nz == 0 || !(resolution > 0.0))
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_alloc: empty grid or resolution <= 0");
  if (frame == DOAPP_FRAME_OG && !(nx == ny && ny == nz))
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_alloc: FRAME_OG grids are cubes");
  CU(cudaSetDevice(ctx->cfg.device));
  const size_t cells = (size_t)nx * ny * nz;
  if (cells >= (1ull << 32))
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_alloc: %zu cells (limit 2^32 - 1)", cells);
  const size_t bytes = cells * (kind == DOAPP_GRID_F32 ? 4 : 1);
  ST(ensure(ctx, ctx->grid_cells, bytes));
  if (kind == DOAPP_GRID_U8) {
    CU(cudaMemsetAsync(ctx->grid_cells.p, 0, bytes, ctx->stream)); // occupancy_grid.cpp:15-16
  } else {
    // +inf == 0x7f800000: fill through a host pattern (distance "nothing nearby")
    std::vector<float> inf(cells, INFINITY);
    CU(cudaMemcpyAsync(ctx->grid_cells.p, inf.data(), bytes, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  GridDev g{};
  g.cells = ctx->grid_cells.p;
  g.kind = kind;
  g.frame = frame;
  g.nx = nx;
  g.ny = ny;
  g.nz = nz;
  g.res = (float)resolution;                                  // cpu_distance_grid.cu:69
  g.x_off = (float)(0.5 * (double)nx * resolution);           // :70-71
  g.y_off = (float)(0.5 * (double)ny * resolution);           // :72-73
  g.z_max = (float)((double)nz * resolution);                 // :81-82
  g.res_d = resolution;
  g.half_d = ((double)nx * resolution) / 2.0;                 // occupancy_grid.cpp:31
  {
    // voxel-index divides (cpu_distance_grid.cu:120-122): every dividend the CDG frame can
    // produce lies in [0, max(2*x_off, 2*y_off, z_max)]; check the truncated quotient of the
    // reciprocal-multiply path against the IEEE divide for ALL of them
    volatile float rr = 1.0f / g.res;
    g.res_rcp = rr;
    float top = fmaxf(fmaxf(2.0f * g.x_off, 2.0f * g.y_off), g.z_max);
    bool ok = false;
    if (frame == DOAPP_FRAME_CDG && top < dm::kFastDivHi)
      ST(verify_fast_div(ctx, g.res, 0u, f2u(top), 1, &ok));
    // One unsigned compare per axis replaces (lo <= c) & (c < hi) & (index < n): the threshold
    // is the smallest dividend a >= 0 with trunc(a / res) >= n (the index is monotone in a).  It
    // is equivalent to the float bounds + index test of cpu_distance_grid.cu:115-122 provided the
    // upper float bound is implied by the index test, i.e. the dividend of c == hi already
    // reaches n; otherwise the grid runs the generic kernel.
    auto threshold = [&](uint32_t n, float top_dividend, uint32_t *bits) -> bool {
      auto idx = [&](uint32_t b) -> double {
        float a;
        memcpy(&a, &b, 4);
        volatile float q = a / g.res; // IEEE RN, as __fdiv_rn
        return std::trunc((double)q);
      };
      if (!(idx(f2u(top_dividend)) >= (double)n)) return false;
      uint32_t lo = 0, hi = f2u(top_dividend); // idx(lo) < n <= idx(hi)
      if (idx(0) >= (double)n) return false;
      while (hi - lo > 1) {
        uint32_t mid = lo + (hi - lo) / 2;
        if (idx(mid) >= (double)n) hi = mid; else lo = mid;
      }
      *bits = hi;
      return true;
    };
    if (ok)
      ok = threshold(nx, 2.0f * g.x_off, &g.tx_bits) && threshold(ny, 2.0f * g.y_off, &g.ty_bits) &&
           threshold(nz, g.z_max, &g.tz_bits) && g.res <= 1.0f;
    g.fast_idx = ok ? 1 : 0;
  }
  ctx->grid = g;
  ctx->grid_bytes = bytes;
  ctx->grid_generation++;
  invalidate_graph(ctx); // captured graphs carry the old GridDev and clear size by value
  return DOAPP_OK;
}

doapp_status doapp_grid_upload(doapp_ctx *ctx, const void *host_cells, size_t num_cells) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!host_cells) return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_upload: null cells");
  if (ctx->grid.kind == DOAPP_GRID_NONE) return fail(ctx, DOAPP_NOT_READY, "no grid allocated");
  size_t cells = (size_t)ctx->grid.nx * ctx->grid.ny * ctx->grid.nz;
  if (num_cells != cells)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_upload: %zu cells, grid has %zu", num_cells,
                cells);
  CU(cudaSetDevice(ctx->cfg.device));
  CU(cudaMemcpyAsync(ctx->grid_cells.p, host_cells, ctx->grid_bytes, cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

doapp_status doapp_grid_download(doapp_ctx *ctx, void *host_cells, size_t num_cells) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!host_cells) return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_download: null cells");
  if (ctx->grid.kind == DOAPP_GRID_NONE) return fail(ctx, DOAPP_NOT_READY, "no grid allocated");
  size_t cells = (size_t)ctx->grid.nx * ctx->grid.ny * ctx->grid.nz;
  if (num_cells != cells)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_download: %zu cells, grid has %zu", num_cells,
                cells);
  CU(cudaSetDevice(ctx->cfg.device));
  CU(cudaMemcpyAsync(host_cells, ctx->grid_cells.p, ctx->grid_bytes, cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

doapp_status doapp_grid_from_msg(doapp_ctx *ctx, double size, double granularity,
                                 const int8_t *grid, size_t len) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!(size > 0.0) || !(granularity > 0.0) || !grid)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_from_msg: bad size/granularity/grid");
  size_t n = (size_t)(size / granularity); // occupancy_grid.cpp:10
  if (n == 0 || n > 2048) return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_from_msg: %zu cells/edge", n);
  if (len != n * n * n)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_from_msg: len %zu != %zu^3", len, n);
  bool same = ctx->grid.kind == DOAPP_GRID_U8 && ctx->grid.frame == DOAPP_FRAME_OG &&
              ctx->grid.nx == n && ctx->grid.res_d == granularity;
  if (!same)
    ST(doapp_grid_alloc(ctx, DOAPP_GRID_U8, DOAPP_FRAME_OG, (uint32_t)n, (uint32_t)n, (uint32_t)n,
                        granularity));
  return doapp_grid_upload(ctx, grid, len);
}

doapp_status doapp_grid_scatter_dev(doapp_ctx *ctx, const float *xyz_dev, size_t num_points,
                                    int8_t val, int clear_first) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (ctx->grid.kind != DOAPP_GRID_U8)
    return fail(ctx, DOAPP_NOT_READY, "doapp_grid_scatter: needs a U8 grid");
  if (num_points && !xyz_dev) return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_scatter: null cloud");
  CU(cudaSetDevice(ctx->cfg.device));
  if (clear_first) CU(cudaMemsetAsync(ctx->grid_cells.p, 0, ctx->grid_bytes, ctx->stream));
  ScatterParams sp{};
  sp.xyz = xyz_dev;
  sp.num_points = num_points;
  sp.grid = ctx->grid;
  sp.cells = ctx->grid_cells.p;
  sp.val = val;
  CU(launch_scatter(sp, ctx->sm_count, ctx->stream));
  if (num_points) ctx->launches += 1;
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

doapp_status doapp_grid_scatter(doapp_ctx *ctx, const float *xyz, size_t num_points, int8_t val,
                                int clear_first) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (num_points && !xyz) return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_scatter: null cloud");
  CU(cudaSetDevice(ctx->cfg.device));
  ST(ensure(ctx, ctx->cloud, num_points * 3));
  if (num_points)
    CU(cudaMemcpyAsync(ctx->cloud.p, xyz, num_points * 3 * sizeof(float), cudaMemcpyHostToDevice,
                       ctx->stream));
  return doapp_grid_scatter_dev(ctx, ctx->cloud.p, num_points, val, clear_first);
}

doapp_status doapp_grid_from_depth(doapp_ctx *ctx, const uint16_t *depth, uint32_t num_cams,
                                   uint32_t width, uint32_t height, const double *intrinsics,
                                   const float *extrinsics, double crop_size, int8_t val,
                                   int clear_first, uint64_t *points_kept) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!depth || !intrinsics || !extrinsics || num_cams == 0 || num_cams > 4 || !(crop_size > 0.0))
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_from_depth: bad argument (1..4 cameras)");
  if (ctx->grid.kind != DOAPP_GRID_U8)
    return fail(ctx, DOAPP_NOT_READY, "doapp_grid_from_depth: needs a U8 grid");
  CU(cudaSetDevice(ctx->cfg.device));
  const size_t pixels = (size_t)num_cams * width * height;
  ST(ensure(ctx, ctx->depth, pixels));
  ST(ensure(ctx, ctx->vbad, 1));
  CU(cudaMemcpyAsync(ctx->depth.p, depth, pixels * sizeof(uint16_t), cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(cudaMemsetAsync(ctx->vbad.p, 0, sizeof(unsigned long long), ctx->stream));
  if (clear_first) CU(cudaMemsetAsync(ctx->grid_cells.p, 0, ctx->grid_bytes, ctx->stream));
  DepthParams p{};
  p.depth = ctx->depth.p;
  p.cams = num_cams;
  p.w = width;
  p.h = height;
  for (uint32_t c = 0; c < num_cams; ++c) {
    const double fx = intrinsics[4 * c], fy = intrinsics[4 * c + 1];
    const double unit_scaling = (double)0.001f; // DepthTraits<uint16_t>::toMeters(1)
    p.cam[c].cx = (float)intrinsics[4 * c + 2];   // mapping_node.cpp:97-98
    p.cam[c].cy = (float)intrinsics[4 * c + 3];
    p.cam[c].kx = (float)(unit_scaling / fx);     // :101-103
    p.cam[c].ky = (float)(unit_scaling / fy);
    for (int i = 0; i < 9; ++i) p.cam[c].r[i] = extrinsics[12 * c + i];
    for (int i = 0; i < 3; ++i) p.cam[c].t[i] = extrinsics[12 * c + 9 + i];
  }
  const float half = (float)(crop_size / 2.0), full = (float)crop_size; // :262-263 (Vector4f)
  p.crop_lo[0] = -half; p.crop_lo[1] = -half; p.crop_lo[2] = 0.0f;
  p.crop_hi[0] = half;  p.crop_hi[1] = half;  p.crop_hi[2] = full;
  p.grid = ctx->grid;
  p.cells = ctx->grid_cells.p;
  p.val = val;
  p.kept = ctx->vbad.p;
  CU(launch_depth_scatter(p, ctx->sm_count, ctx->stream));
  ctx->launches += 1;
  ctx->grid_generation++;
  unsigned long long kept = 0;
  CU(cudaMemcpyAsync(&kept, ctx->vbad.p, sizeof(kept), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (points_kept) *points_kept = kept;
  return DOAPP_OK;
}

doapp_status doapp_grid_distance_field(doapp_ctx *ctx) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (ctx->grid.kind != DOAPP_GRID_U8)
    return fail(ctx, DOAPP_NOT_READY, "doapp_grid_distance_field: needs a U8 grid");
  const GridDev &g0 = ctx->grid;
  if (g0.nx > kEdtMaxDim || g0.ny > kEdtMaxDim || g0.nz > kEdtMaxDim)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_distance_field: dimensions above %u", kEdtMaxDim);
  CU(cudaSetDevice(ctx->cfg.device));
  const size_t cells = (size_t)g0.nx * g0.ny * g0.nz;
  DevBuf<int32_t> ta, tb;
  DevBuf<uint8_t> dist;
  doapp_status st = [&]() -> doapp_status {
    ST(ensure(ctx, ta, cells));
    ST(ensure(ctx, tb, cells));
    ST(ensure(ctx, dist, cells * sizeof(float)));
    CU(launch_distance_field(ctx->grid_cells.p, g0.nx, g0.ny, g0.nz, g0.res, ta.p, tb.p,
                             (float *)dist.p, ctx->stream));
    ctx->launches += 3;
    CU(cudaStreamSynchronize(ctx->stream));
    return DOAPP_OK;
  }();
  release(ta);
  release(tb);
  if (st != DOAPP_OK) {
    release(dist);
    return st;
  }
  std::swap(ctx->grid_cells, dist); // the distance field replaces the occupancy cells
  release(dist);
  ctx->grid.cells = ctx->grid_cells.p;
  ctx->grid.kind = DOAPP_GRID_F32;
  ctx->grid_bytes = cells * sizeof(float);
  ctx->grid_generation++;
  invalidate_graph(ctx);
  return DOAPP_OK;
}

// ---- distance field with NO allocation (no
// counterpart in the reference).  The two int32 scratch planes of the distance transform live
// in a file-scope device array instead of cudaMalloc'd buffers: every GPU that loads the
// library holds DOAPP_NOALLOC_MAX_CELLS * 8 bytes (16 MiB) until the process exits.  In host
// code `g_edt_scratch` is only a stand-in for the symbol; the device address is looked up once
// per device with cudaGetSymbolAddress (that device current) and handed to the kernels.
} // extern "C"
static __device__ __align__(256) int32_t g_edt_scratch[2 * DOAPP_NOALLOC_MAX_CELLS];
namespace {
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
int32_t *g_noalloc_ptr[kNoallocMaxDevices] = {};
} // namespace
extern "C" {

doapp_status doapp_distance_field_noalloc(const uint8_t *occ_dev, uint32_t nx, uint32_t ny,
                                          uint32_t nz, double resolution, float *dist_dev,
                                          void *stream) {
  doapp_ctx *ctx = nullptr; // errors go to the thread-local slot read by doapp_last_error(NULL)
  if (!occ_dev || !dist_dev)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_distance_field_noalloc: null device pointer");
  if (nx == 0 || ny == 0 || nz == 0 || !(resolution > 0.0))
    return fail(nullptr, DOAPP_INVALID_ARG,
                "doapp_distance_field_noalloc: empty grid or resolution <= 0");
  if (nx > kEdtMaxDim || ny > kEdtMaxDim || nz > kEdtMaxDim)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_distance_field_noalloc: dimensions above %u",
                kEdtMaxDim);
  const uint64_t cells = (uint64_t)nx * ny * nz;
  if (cells > (uint64_t)DOAPP_NOALLOC_MAX_CELLS)
    return fail(nullptr, DOAPP_INVALID_ARG,
                "doapp_distance_field_noalloc: %llu cells exceed the static scratch of %llu cells "
                "(DOAPP_NOALLOC_MAX_CELLS); use doapp_grid_distance_field for larger grids",
                (unsigned long long)cells, (unsigned long long)DOAPP_NOALLOC_MAX_CELLS);
  int dev = -1;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoallocMaxDevices)
    return fail(nullptr, DOAPP_INVALID_ARG, "doapp_distance_field_noalloc: device %d", dev);
  int32_t *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) { // once per device, with that device current
      void *p = nullptr;
      CU(cudaGetSymbolAddress(&p, g_edt_scratch));
      g_noalloc_ptr[dev] = static_cast<int32_t *>(p);
    }
    scratch = g_noalloc_ptr[dev];
  }
  // rows -> tmp_a, lines along y -> tmp_b, lines along z -> dist: every cell of both planes is
  // written before it is read, so nothing has to start at zero
  CU(launch_distance_field(occ_dev, nx, ny, nz, (float)resolution, scratch,
                           scratch + DOAPP_NOALLOC_MAX_CELLS, dist_dev, (cudaStream_t)stream));
  return DOAPP_OK;
}

doapp_status doapp_grid_lookup(doapp_ctx *ctx, const float *xyz, size_t num_points,
                               int64_t *cell_index, uint8_t *collides) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (ctx->grid.kind == DOAPP_GRID_NONE) return fail(ctx, DOAPP_NOT_READY, "no grid allocated");
  if (num_points == 0) return DOAPP_OK;
  if (!xyz || !cell_index || !collides)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_grid_lookup: null argument");
  CU(cudaSetDevice(ctx->cfg.device));
  ST(ensure(ctx, ctx->pts, num_points * 3));
  ST(ensure(ctx, ctx->lookup_idx, num_points));
  ST(ensure(ctx, ctx->lookup_coll, num_points));
  CU(cudaMemcpyAsync(ctx->pts.p, xyz, num_points * 3 * sizeof(float), cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(launch_lookup(ctx->grid, ctx->pts.p, num_points, ctx->cfg.collision_distance,
                   ctx->lookup_idx.p, ctx->lookup_coll.p, ctx->stream));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(cell_index, ctx->lookup_idx.p, num_points * sizeof(long long),
                     cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(collides, ctx->lookup_coll.p, num_points, cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

doapp_status doapp_set_limits(doapp_ctx *ctx, const float *min_d, const float *max_d) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!min_d || !max_d) {
    ctx->use_limits = 0;
  } else {
    for (uint32_t j = 0; j < ctx->cfg.waypoint_dim; ++j) {
      ctx->lim_min[j] = min_d[j];
      ctx->lim_max[j] = max_d[j];
    }
    ctx->use_limits = 1;
  }
  invalidate_graph(ctx);
  return DOAPP_OK;
}

// -------------------------------------------------------------------- FK model

void doapp_fk_model_rexarm(doapp_fk_model *m) {
  if (!m) return;
  memset(m, 0, sizeof(*m));
  m->struct_size = sizeof(*m);
  const float origin[5][3] = {{0.0f, -0.0015f, 0.0715f}, // rexarm.urdf:64
                              {0.0f, 0.0f, 0.045f},      // :83
                              {0.0f, 0.0f, 0.1f},        // :109
                              {0.0f, 0.0f, 0.1f},        // :128
                              {0.0f, -0.0125f, 0.08f}};  // :147
  const int32_t axis[5] = {2, 1, 1, 1, 2};               // :65,84,110,129,148
  memcpy(m->joint_origin, origin, sizeof(origin));
  memcpy(m->joint_axis, axis, sizeof(axis));
  // every link box (visual origin o, size b) gets two spheres on its long side at o -+ long/4
  // with radius max(short sides)/2; the fingers box likewise; one sphere at gripper_center
  struct Box { int link; double o[3]; double b[3]; };
  const Box boxes[6] = {{0, {0.0, 0.0, 0.031}, {0.035, 0.03, 0.05}},        // :52,55
                        {1, {0.0, 0.0, 0.086}, {0.035, 0.03, 0.05}},        // :71,74
                        {2, {0.0, 0.0, 0.08575}, {0.032, 0.031, 0.05}},     // :97,100
                        {3, {0.0, -0.00175, 0.0645}, {0.032, 0.05, 0.031}}, // :116,119
                        {4, {0.0, 0.00125, 0.019}, {0.031, 0.05, 0.031}},   // :135,138
                        {4, {0.0, 0.01, 0.07}, {0.0275, 0.01, 0.07}}};      // :180,189
  uint32_t q = 0;
  for (const Box &bx : boxes) {
    int lg = 0;
    for (int i = 1; i < 3; ++i)
      if (bx.b[i] > bx.b[lg]) lg = i;
    double rad = 0.0;
    for (int i = 0; i < 3; ++i)
      if (i != lg && bx.b[i] > rad) rad = bx.b[i];
    for (int sgn = -1; sgn <= 1; sgn += 2) {
      doapp_fk_sphere &sp = m->spheres[q++];
      sp.link = bx.link;
      for (int i = 0; i < 3; ++i) sp.centre[i] = (float)bx.o[i];
      sp.centre[lg] = (float)(bx.o[lg] + sgn * 0.25 * bx.b[lg]);
      sp.radius = (float)(0.5 * rad);
    }
  }
  m->spheres[q++] = doapp_fk_sphere{4, {0.0f, -0.005f, 0.07f}, 0.01f}; // :197
  m->num_spheres = q;
}

doapp_status doapp_set_fk_model(doapp_ctx *ctx, const doapp_fk_model *m) {
  if (!ctx) return DOAPP_INVALID_ARG;
  CU(cudaSetDevice(ctx->cfg.device));
  CU(cudaStreamSynchronize(ctx->stream));
  invalidate_graph(ctx);
  if (!m) {
    ctx->fk_on = false;
    return DOAPP_OK;
  }
  if (m->struct_size != sizeof(doapp_fk_model))
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_set_fk_model: struct_size %u != %zu", m->struct_size,
                sizeof(doapp_fk_model));
  if (m->num_spheres == 0 || m->num_spheres > DOAPP_FK_MAX_SPHERES)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_set_fk_model: num_spheres %u not in 1..%d",
                m->num_spheres, DOAPP_FK_MAX_SPHERES);
  FkModelDev h{};
  for (int l = 0; l < 5; ++l) {
    if (m->joint_axis[l] < 0 || m->joint_axis[l] > 2)
      return fail(ctx, DOAPP_INVALID_ARG, "doapp_set_fk_model: joint_axis[%d] = %d", l, m->joint_axis[l]);
    h.axis[l] = m->joint_axis[l];
    for (int i = 0; i < 3; ++i) h.origin[l][i] = m->joint_origin[l][i];
  }
  uint32_t q = 0;
  for (int l = 0; l < 5; ++l) { // stable sort by link
    h.first[l] = q;
    for (uint32_t i = 0; i < m->num_spheres; ++i) {
      const doapp_fk_sphere &sp = m->spheres[i];
      if (sp.link < 0 || sp.link > 4 || !(sp.radius >= 0.0f))
        return fail(ctx, DOAPP_INVALID_ARG, "doapp_set_fk_model: sphere %u link %d radius %g", i,
                    sp.link, (double)sp.radius);
      if (sp.link != l) continue;
      h.order[q] = i;
      for (int a = 0; a < 3; ++a) h.centre[q][a] = sp.centre[a];
      volatile float thr = sp.radius + ctx->cfg.collision_distance; // one fp32 add, as the oracle
      h.thr[q] = thr;
      ++q;
    }
  }
  h.first[5] = q;
  ST(ensure(ctx, ctx->fk_dev, 1));
  CU(cudaMemcpyAsync(ctx->fk_dev.p, &h, sizeof(h), cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->fk_on = true;
  ctx->fk_host = h;
  ctx->fk_spheres = m->num_spheres;
  return DOAPP_OK;
}

// ----------------------------------------------------------------------- noise

doapp_status doapp_noise_generate(doapp_ctx *ctx, uint32_t iteration, float *noise_out) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!noise_out) return fail(ctx, DOAPP_INVALID_ARG, "doapp_noise_generate: null output");
  CU(cudaSetDevice(ctx->cfg.device));
  const doapp_config &c = ctx->cfg;
  size_t total = (size_t)ctx->S * c.noise_waypoints * c.waypoint_dim;
  ST(ensure(ctx, ctx->dbg_a, total));
  CostParams p = base_cost_params(ctx);
  p.iteration = iteration;
  CU(launch_noise_fill(p, c.waypoint_dim, ctx->dbg_a.p, ctx->stream));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(noise_out, ctx->dbg_a.p, total * sizeof(float), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

// ------------------------------------------------------------- cost evaluation

doapp_status doapp_stage_inputs(doapp_ctx *ctx, const float *base_traj, const float *start,
                                const float *noise) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!base_traj || !start)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_stage_inputs: null trajectory or start");
  CU(cudaSetDevice(ctx->cfg.device));
  const doapp_config &c = ctx->cfg;
  ST(upload_traj(ctx, base_traj, false));
  ST(upload_start(ctx, start));
  if (noise) {
    size_t total = (size_t)ctx->S * c.noise_waypoints * c.waypoint_dim;
    ST(ensure(ctx, ctx->noise, total));
    CU(cudaMemcpyAsync(ctx->noise.p, noise, total * sizeof(float), cudaMemcpyHostToDevice,
                       ctx->stream));
  }
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

static doapp_status enqueue_cost_eval(doapp_ctx *ctx, uint32_t iteration, int noise_mode) {
  const doapp_config &c = ctx->cfg;
  CU(cudaMemsetAsync(ctx->keys_float.p, 0xFF, sizeof(unsigned long long) * c.num_climbers,
                     ctx->stream));
  CostParams p = base_cost_params(ctx);
  p.noise_mode = noise_mode;
  p.iteration = iteration;
  if (noise_mode == NOISE_EXTERNAL) {
    p.noise = ctx->noise.p;
    p.noise_m = c.samples_per_climber;
    p.noise_t0 = c.sample_offset;
  }
  p.scores = ctx->scores.p;
  p.collisions = ctx->coll.p;
  p.keys_float = ctx->keys_float.p;
  CU(launch_cost(p, c.waypoint_dim, ctx->sm_count, ctx->stream));
  ctx->launches += 1;
  return DOAPP_OK;
}

doapp_status doapp_read_results(doapp_ctx *ctx, float *scores, int32_t *collisions,
                                int64_t *argmin, float *min_score) {
  if (!ctx) return DOAPP_INVALID_ARG;
  CU(cudaSetDevice(ctx->cfg.device));
  const doapp_config &c = ctx->cfg;
  std::vector<unsigned long long> keys(c.num_climbers);
  if (scores)
    CU(cudaMemcpyAsync(scores, ctx->scores.p, ctx->S * sizeof(float), cudaMemcpyDeviceToHost,
                       ctx->stream));
  if (collisions)
    CU(cudaMemcpyAsync(collisions, ctx->coll.p, ctx->S * sizeof(int32_t), cudaMemcpyDeviceToHost,
                       ctx->stream));
  CU(cudaMemcpyAsync(keys.data(), ctx->keys_float.p, keys.size() * sizeof(unsigned long long),
                     cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  // planning_state.cu:113-122 over the per-climber winners: strict <, first wins
  float lowest = INFINITY;
  int64_t best = -1;
  for (uint32_t cl = 0; cl < c.num_climbers; ++cl) {
    uint32_t hi = (uint32_t)(keys[cl] >> 32);
    float sc;
    memcpy(&sc, &hi, 4);
    if (hi == 0xFFFFFFFFu) continue;
    if (sc < lowest) {
      lowest = sc;
      best = (int64_t)((uint64_t)cl * ctx->m_global + (uint32_t)keys[cl]);
    }
  }
  if (argmin) *argmin = best;
  if (min_score) *min_score = lowest;
  return DOAPP_OK;
}

doapp_status doapp_cost_eval(doapp_ctx *ctx, const float *base_traj, const float *start,
                             const float *noise, uint32_t iteration, float *scores,
                             int32_t *collisions, int64_t *argmin, float *min_score) {
  if (!ctx) return DOAPP_INVALID_ARG;
  ST(doapp_stage_inputs(ctx, base_traj, start, noise));
  ST(enqueue_cost_eval(ctx, iteration, noise ? NOISE_EXTERNAL : NOISE_PHILOX));
  return doapp_read_results(ctx, scores, collisions, argmin, min_score);
}

doapp_status doapp_cost_eval_dev(doapp_ctx *ctx, uint32_t iteration, int use_staged_noise,
                                 uint32_t repeats, float *elapsed_ms) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (use_staged_noise && !ctx->noise.p)
    return fail(ctx, DOAPP_NOT_READY, "doapp_cost_eval_dev: no staged noise");
  CU(cudaSetDevice(ctx->cfg.device));
  if (repeats == 0) repeats = 1;
  CU(cudaEventRecord(ctx->ev0, ctx->stream));
  for (uint32_t r = 0; r < repeats; ++r)
    ST(enqueue_cost_eval(ctx, iteration, use_staged_noise ? NOISE_EXTERNAL : NOISE_PHILOX));
  CU(cudaEventRecord(ctx->ev1, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  float ms = 0.0f;
  CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  if (elapsed_ms) *elapsed_ms = ms;
  return DOAPP_OK;
}

// --------------------------------------------------------------------- planner

doapp_status doapp_plan_lerp(const float *start, const float *goal, uint32_t n, uint32_t d,
                             float *out) {
  if (!start || !goal || !out || n == 0 || d == 0) return DOAPP_INVALID_ARG;
  // planning_state.cu:50-52: step = (goal - current) / num_waypoints
  for (uint32_t j = 0; j < d; ++j) {
    volatile float step = (goal[j] - start[j]) / (float)n;
    for (uint32_t w = 0; w < n; ++w) {
      volatile float inc = (float)(w + 1) * step;
      out[(size_t)w * d + j] = start[j] + inc;
    }
  }
  return DOAPP_OK;
}

doapp_status doapp_plan_async(doapp_ctx *ctx, const float *start, const float *goal,
                              const float *init_traj, uint32_t iters, const float *noise) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!start || !init_traj)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_plan: null start or init_traj");
  (void)goal; // only the message assembly needs it (host side)
  CU(cudaSetDevice(ctx->cfg.device));
  const doapp_config &c = ctx->cfg;
  ST(upload_traj(ctx, init_traj, true));
  ST(upload_start(ctx, start));
  int mode = NOISE_PHILOX;
  if (noise) {
    size_t total = (size_t)iters * c.num_climbers * ctx->m_global * c.noise_waypoints *
                   c.waypoint_dim;
    ST(ensure(ctx, ctx->noise, total));
    CU(cudaMemcpyAsync(ctx->noise.p, noise, total * sizeof(float), cudaMemcpyHostToDevice,
                       ctx->stream));
    mode = NOISE_EXTERNAL;
  }
  ctx->plan_staged = true;
  ctx->plan_noise_mode = mode;
  ctx->plan_iters = iters;
  ST(run_plan(ctx, iters, mode));
  ctx->plan_pending = true;
  return DOAPP_OK;
}

doapp_status doapp_plan_wait(doapp_ctx *ctx, float *best_traj, float *best_vel, float *best_score,
                             int64_t *best_climber) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!ctx->plan_pending) return fail(ctx, DOAPP_NOT_READY, "doapp_plan_wait: no plan in flight");
  CU(cudaSetDevice(ctx->cfg.device));
  const size_t nd = (size_t)ctx->cfg.num_waypoints * ctx->cfg.waypoint_dim;
  CU(cudaMemcpyAsync(ctx->h_out.data(), ctx->out.p, (4 + 2 * nd) * sizeof(float),
                     cudaMemcpyDeviceToHost, ctx->stream));
  unsigned int p2_err = 0, xr_err = 0;
  if (ctx->p2_ran)
    CU(cudaMemcpyAsync(&p2_err, ctx->p2_error.p, sizeof(p2_err), cudaMemcpyDeviceToHost, ctx->stream));
  if (ctx->xr_on)
    CU(cudaMemcpyAsync(&xr_err, ctx->xr_error.p, sizeof(xr_err), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->plan_pending = false;
  if (xr_err) {
    CU(cudaMemsetAsync(ctx->xr_error.p, 0, sizeof(unsigned int), ctx->stream));
    return fail(ctx, DOAPP_CUDA_ERROR,
                "planner: a rank's argmin key did not arrive in the peer mailbox (timed out)");
  }
  if (p2_err) {
    CU(cudaMemsetAsync(ctx->p2_error.p, 0, sizeof(unsigned int), ctx->stream));
    return fail(ctx, DOAPP_CUDA_ERROR,
                "persistent planner: the argmin key exchange timed out (a rank did not arrive)");
  }
  if (best_score) *best_score = ctx->h_out[0];
  if (best_climber) memcpy(best_climber, &ctx->h_out[2], 8);
  if (best_traj) memcpy(best_traj, &ctx->h_out[4], nd * sizeof(float));
  if (best_vel) memcpy(best_vel, &ctx->h_out[4 + nd], nd * sizeof(float));
  return DOAPP_OK;
}

doapp_status doapp_plan(doapp_ctx *ctx, const float *start, const float *goal,
                        const float *init_traj, uint32_t iters, const float *noise,
                        float *best_traj, float *best_vel, float *best_score,
                        int64_t *best_climber) {
  ST(doapp_plan_async(ctx, start, goal, init_traj, iters, noise));
  return doapp_plan_wait(ctx, best_traj, best_vel, best_score, best_climber);
}

doapp_status doapp_plan_state(doapp_ctx *ctx, float *trajectories, float *scores,
                              int32_t *rounded, uint8_t *found_better) {
  if (!ctx) return DOAPP_INVALID_ARG;
  CU(cudaSetDevice(ctx->cfg.device));
  const doapp_config &c = ctx->cfg;
  const size_t k = c.num_climbers, n = c.num_waypoints, d = c.waypoint_dim, npad = ctx->npad;
  CU(cudaStreamSynchronize(ctx->stream));
  if (trajectories) {
    std::vector<float> soa(k * d * npad);
    CU(cudaMemcpy(soa.data(), traj_half_ptr(ctx, ctx->traj_s), soa.size() * sizeof(float),
                  cudaMemcpyDeviceToHost));
    for (size_t cl = 0; cl < k; ++cl)
      for (size_t w = 0; w < n; ++w)
        for (size_t j = 0; j < d; ++j)
          trajectories[(cl * n + w) * d + j] = soa[(cl * d + j) * npad + w];
  }
  if (scores || rounded) {
    std::vector<float> sc(k);
    CU(cudaMemcpy(sc.data(), ctx->final_scores.p, k * sizeof(float), cudaMemcpyDeviceToHost));
    for (size_t cl = 0; cl < k; ++cl) {
      if (scores) scores[cl] = sc[cl];
      if (rounded)
        rounded[cl] = (sc[cl] < 2147483648.0f) ? (int32_t)ceilf(sc[cl]) : 0x7FFFFFFF;
    }
  }
  if (found_better)
    CU(cudaMemcpy(found_better, ctx->found.p, k, cudaMemcpyDeviceToHost));
  return DOAPP_OK;
}

doapp_status doapp_plan_bench(doapp_ctx *ctx, uint32_t iters, uint32_t repeats,
                              float *elapsed_ms) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!ctx->plan_staged) return fail(ctx, DOAPP_NOT_READY, "doapp_plan_bench: call doapp_plan first");
  if (ctx->plan_noise_mode == NOISE_EXTERNAL && iters > ctx->plan_iters)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_plan_bench: staged noise covers %u iterations",
                ctx->plan_iters);
  CU(cudaSetDevice(ctx->cfg.device));
  if (repeats == 0) repeats = 1;
  // make sure the graph exists before the timed region (no untimed pre-run
  // otherwise: the caller may have flushed L2 on purpose)
  const float *nptr = ctx->plan_noise_mode == NOISE_EXTERNAL ? ctx->noise.p : nullptr;
  const bool have_graph = ctx->graph_exec && ctx->graph_iters == iters &&
                          ctx->graph_noise_mode == ctx->plan_noise_mode &&
                          ctx->graph_noise_ptr == nptr &&
                          ctx->graph_grid_generation == ctx->grid_generation;
  if (ctx->use_graph && !have_graph) {
    ST(reset_climbers(ctx));
    ST(run_plan(ctx, iters, ctx->plan_noise_mode));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  CU(cudaEventRecord(ctx->ev0, ctx->stream));
  for (uint32_t r = 0; r < repeats; ++r) {
    ST(reset_climbers(ctx));
    ST(run_plan(ctx, iters, ctx->plan_noise_mode));
  }
  CU(cudaEventRecord(ctx->ev1, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  float ms = 0.0f;
  CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  if (elapsed_ms) *elapsed_ms = ms;
  ctx->plan_pending = true;
  return DOAPP_OK;
}

doapp_status doapp_replan_dynamic(doapp_ctx *ctx, const float *start, const float *goal,
                                  const float *init_traj, uint32_t iters, const float *clouds,
                                  size_t num_points, float *best_traj, float *best_score,
                                  float *elapsed_ms) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!start || !init_traj) return fail(ctx, DOAPP_INVALID_ARG, "doapp_replan_dynamic: null argument");
  if (ctx->grid.kind != DOAPP_GRID_U8)
    return fail(ctx, DOAPP_NOT_READY, "doapp_replan_dynamic: needs a U8 grid");
  (void)goal;
  CU(cudaSetDevice(ctx->cfg.device));
  const doapp_config &c = ctx->cfg;
  ST(upload_traj(ctx, init_traj, true));
  ST(upload_start(ctx, start));
  const size_t cloud_elems = (size_t)iters * num_points * 3;
  if (!clouds && cloud_elems && ctx->cloud_staged != cloud_elems)
    return fail(ctx, DOAPP_NOT_READY,
                "doapp_replan_dynamic: clouds == NULL reuses the clouds of the previous call, "
                "which staged %zu floats, not %zu", ctx->cloud_staged, cloud_elems);
  ST(ensure(ctx, ctx->cloud, cloud_elems));
  if (!ctx->copy_stream) CU(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  while (ctx->cloud_events.size() < iters) {
    cudaEvent_t e = nullptr;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ctx->cloud_events.push_back(e);
  }
  CU(cudaStreamSynchronize(ctx->stream)); // inputs staged; the previous loop no longer reads the clouds
  if (clouds && cloud_elems) {
    // one copy + one event per cloud on the copy stream: iteration it waits for cloud it only
    for (uint32_t it = 0; it < iters; ++it) {
      const size_t off = (size_t)it * num_points * 3;
      CU(cudaMemcpyAsync(ctx->cloud.p + off, clouds + off, num_points * 3 * sizeof(float),
                         cudaMemcpyHostToDevice, ctx->copy_stream));
      CU(cudaEventRecord(ctx->cloud_events[it], ctx->copy_stream));
    }
    ctx->cloud_staged = cloud_elems;
    if (!ctx->replan_ext_wait_ok) CU(cudaStreamSynchronize(ctx->copy_stream));
  }
  CU(cudaEventRecord(ctx->ev0, ctx->stream));
  bool capturing = false;
  const size_t k = c.num_climbers;
  // the whole loop (per iteration: clear + scatter + re-score + cost pass) as one sequence;
  // captured once into a CUDA graph and replayed while shapes and buffers stay the same
  auto enqueue = [&](uint64_t *kernels) -> doapp_status {
    uint64_t nk = 0;
    reset_plan_state(ctx);
    CU(cudaMemsetAsync(ctx->keys.p, 0xFF, sizeof(unsigned long long) * 3 * k, ctx->stream));
    for (uint32_t it = 0; it < iters; ++it) {
      CU(cudaMemsetAsync(ctx->grid_cells.p, 0, ctx->grid_bytes, ctx->stream));
      if (ctx->replan_ext_wait_ok && num_points)
        CU(cudaStreamWaitEvent(ctx->stream, ctx->cloud_events[it],
                               capturing ? cudaEventWaitExternal : cudaEventWaitDefault));
      ScatterParams sp{};
      sp.xyz = ctx->cloud.p + (size_t)it * num_points * 3;
      sp.num_points = num_points;
      sp.grid = ctx->grid;
      sp.cells = ctx->grid_cells.p;
      sp.val = 1;
      CU(launch_scatter(sp, ctx->sm_count, ctx->stream));
      if (num_points) nk += 1;
      // the world moved: apply the previous iteration's keep-best, then re-score the kept
      // trajectories against the new grid (their key seeds cur_hi for this iteration)
      ST(launch_pass(ctx, climber_cost_params(ctx), it ? PEND_KEEP_BEST : PEND_NONE,
                     it ? it - 1 : 0, NOISE_PHILOX, nullptr, true));
      CostParams p = base_cost_params(ctx);
      p.noise_mode = NOISE_PHILOX;
      p.iteration = it;
      ST(launch_pass(ctx, p, PEND_SEED, 0, NOISE_PHILOX, nullptr, true));
      nk += 2;
      if (ctx->comm) { // sharded samples: every rank needs the global winner
        NcclApi *nccl = nccl_api(nullptr);
        unsigned long long *kb = ctx->keys.p + (size_t)ctx->key_w * k;
        ncclResult_t r = nccl->AllReduce(kb, kb, k, ncclUint64, ncclMin, ctx->comm, ctx->stream);
        if (r != ncclSuccess)
          return fail(ctx, DOAPP_NCCL_ERROR, "ncclAllReduce: %s",
                      nccl->GetErrorString ? nccl->GetErrorString(r) : "error");
      }
    }
    CostParams p = climber_cost_params(ctx);
    p.scores = ctx->final_scores.p;
    p.collisions = ctx->final_coll.p;
    if (iters)
      ST(launch_pass(ctx, p, PEND_KEEP_BEST, iters - 1, NOISE_PHILOX, nullptr, false));
    else
      ST(launch_pass(ctx, p, PEND_NONE, 0, NOISE_ZERO, nullptr, false));
    FinalParams f{};
    f.traj = traj_half_ptr(ctx, ctx->traj_s);
    f.start = ctx->start.p;
    f.scores = ctx->final_scores.p;
    f.found_better = ctx->found.p;
    f.k = c.num_climbers;
    f.n = c.num_waypoints;
    f.npad = ctx->npad;
    f.d = c.waypoint_dim;
    f.dt = c.delta_t;
    f.out = ctx->out.p;
    CU(launch_final(f, ctx->stream));
    nk += 2;
    *kernels = nk;
    return DOAPP_OK;
  };
  if (!ctx->use_graph) {
    uint64_t nk = 0;
    ST(enqueue(&nk));
    ctx->launches += nk;
  } else {
    const bool hit = ctx->replan_exec && ctx->replan_iters == iters &&
                     ctx->replan_points == num_points && ctx->replan_cloud == ctx->cloud.p &&
                     ctx->replan_cells == ctx->grid_cells.p &&
                     ctx->replan_grid_generation == ctx->grid_generation &&
                     ctx->replan_waits == ctx->replan_ext_wait_ok;
    if (!hit) {
      if (ctx->replan_exec) cudaGraphExecDestroy(ctx->replan_exec);
      ctx->replan_exec = nullptr;
      uint64_t nk = 0;
      for (int attempt = 0; attempt < 2 && !ctx->replan_exec; ++attempt) {
        cudaGraph_t graph = nullptr;
        CU(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
        capturing = true;
        doapp_status st = enqueue(&nk);
        capturing = false;
        cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
        if (st == DOAPP_OK && ce == cudaSuccess) {
          ce = cudaGraphInstantiate(&ctx->replan_exec, graph, 0);
          if (ce != cudaSuccess) ctx->replan_exec = nullptr;
        }
        if (graph) cudaGraphDestroy(graph);
        if (ctx->replan_exec) break;
        if (attempt == 0 && ctx->replan_ext_wait_ok) {
          // external event waits not capturable here: wait for the copies on the host instead
          cudaGetLastError();
          ctx->replan_ext_wait_ok = false;
          CU(cudaStreamSynchronize(ctx->copy_stream));
          continue;
        }
        if (st != DOAPP_OK) return st;
        return fail(ctx, DOAPP_CUDA_ERROR, "replan graph: %s", cudaGetErrorString(ce));
      }
      ctx->replan_waits = ctx->replan_ext_wait_ok;
      ctx->replan_iters = iters;
      ctx->replan_points = num_points;
      ctx->replan_cloud = ctx->cloud.p;
      ctx->replan_cells = ctx->grid_cells.p;
      ctx->replan_grid_generation = ctx->grid_generation;
      ctx->replan_kernels = nk;
      ctx->replan_end[0] = ctx->traj_s;
      ctx->replan_end[1] = ctx->cur_s;
      ctx->replan_end[2] = ctx->key_w;
    }
    CU(cudaGraphLaunch(ctx->replan_exec, ctx->stream));
    ctx->traj_s = ctx->replan_end[0];
    ctx->cur_s = ctx->replan_end[1];
    ctx->key_w = ctx->replan_end[2];
    ctx->launches += ctx->replan_kernels;
  }
  CU(cudaEventRecord(ctx->ev1, ctx->stream));
  ctx->plan_pending = true;
  ST(doapp_plan_wait(ctx, best_traj, nullptr, best_score, nullptr));
  float ms = 0.0f;
  CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  if (elapsed_ms) *elapsed_ms = ms;
  return DOAPP_OK;
}

doapp_status doapp_plan_profile(doapp_ctx *ctx, uint32_t iters, float *cost_ms_total,
                                uint32_t *cost_launches, float *loop_ms) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!ctx->plan_staged)
    return fail(ctx, DOAPP_NOT_READY, "doapp_plan_profile: call doapp_plan first");
  if (ctx->plan_noise_mode == NOISE_EXTERNAL && iters > ctx->plan_iters)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_plan_profile: staged noise covers %u iterations",
                ctx->plan_iters);
  CU(cudaSetDevice(ctx->cfg.device));
  while (ctx->prof_events.size() < 2 * (size_t)(iters ? iters : 1)) {
    cudaEvent_t e = nullptr;
    CU(cudaEventCreate(&e));
    ctx->prof_events.push_back(e);
  }
  ST(reset_climbers(ctx));
  CU(cudaEventRecord(ctx->ev0, ctx->stream));
  ctx->profiling = true;
  doapp_status st = run_plan(ctx, iters, ctx->plan_noise_mode);
  ctx->profiling = false;
  ST(st);
  CU(cudaEventRecord(ctx->ev1, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  float total = 0.0f;
  const uint32_t timed = ctx->prof_launches; // iters launches, or 1 persistent launch
  for (uint32_t it = 0; it < timed; ++it) {
    float ms = 0.0f;
    CU(cudaEventElapsedTime(&ms, ctx->prof_events[2 * it], ctx->prof_events[2 * it + 1]));
    total += ms;
  }
  if (cost_ms_total) *cost_ms_total = total;
  if (cost_launches) *cost_launches = timed;
  if (loop_ms) CU(cudaEventElapsedTime(loop_ms, ctx->ev0, ctx->ev1));
  ctx->plan_pending = true;
  return DOAPP_OK;
}

doapp_status doapp_probe_gather(doapp_ctx *ctx, size_t array_bytes, uint32_t cell_bytes,
                                double *gathers_per_sec) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!gathers_per_sec || array_bytes < 4096 || (cell_bytes != 1 && cell_bytes != 4))
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_probe_gather: bad argument");
  CU(cudaSetDevice(ctx->cfg.device));
  DevBuf<uint8_t> arr;
  DevBuf<uint32_t> sink;
  doapp_status st = [&]() -> doapp_status {
    ST(ensure(ctx, arr, array_bytes));
    ST(ensure(ctx, sink, 1));
    CU(cudaMemsetAsync(arr.p, 1, array_bytes, ctx->stream));
    const uint32_t blocks = (uint32_t)ctx->sm_count * 8, per_thread = 1024;
    const uint64_t cells = array_bytes / cell_bytes;
    for (int rep = 0; rep < 3; ++rep) { // the last repetition is the measurement
      CU(cudaEventRecord(ctx->ev0, ctx->stream));
      CU(launch_gather_probe(arr.p, cells, cell_bytes, blocks, per_thread, sink.p, ctx->stream));
      CU(cudaEventRecord(ctx->ev1, ctx->stream));
      ctx->launches += 1;
    }
    CU(cudaStreamSynchronize(ctx->stream));
    float ms = 0.0f;
    CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    *gathers_per_sec = (double)blocks * 256.0 * per_thread / (ms * 1e-3);
    return DOAPP_OK;
  }();
  release(arr);
  release(sink);
  return st;
}

// ------------------------------------------------------------------- multi-GPU

doapp_status doapp_peer_export(doapp_ctx *ctx, uint8_t handle[DOAPP_PEER_HANDLE_BYTES]) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!handle) return fail(ctx, DOAPP_INVALID_ARG, "doapp_peer_export: null handle");
  static_assert(sizeof(cudaIpcMemHandle_t) <= DOAPP_PEER_HANDLE_BYTES, "IPC handle size");
  CU(cudaSetDevice(ctx->cfg.device));
  // two alternating sets of [ring][k][8 ranks] words (kP2Ring covers the first-generation
  // kernel's three-way rotation as well)
  const size_t k = ctx->cfg.num_climbers, elems = 2 * (size_t)kP2Ring * k * 8;
  ST(ensure(ctx, ctx->mailbox, elems));
  ST(ensure(ctx, ctx->presult, 3 * k));
  CU(cudaMemsetAsync(ctx->mailbox.p, 0xFF, elems * sizeof(unsigned long long), ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, ctx->mailbox.p));
  memset(handle, 0, DOAPP_PEER_HANDLE_BYTES);
  memcpy(handle, &h, sizeof(h));
  return DOAPP_OK;
}

doapp_status doapp_peer_import(doapp_ctx *ctx, const uint8_t *handles, int32_t rank,
                               int32_t nranks) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!handles && nranks == 0) { // disable: back to the NCCL merge
    ctx->peer_ready = false;
    return DOAPP_OK;
  }
  if (!handles || nranks < 1 || nranks > 8 || rank < 0 || rank >= nranks)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_peer_import: bad argument (1..8 ranks)");
  if (!ctx->mailbox.p) return fail(ctx, DOAPP_NOT_READY, "doapp_peer_import: call doapp_peer_export first");
  CU(cudaSetDevice(ctx->cfg.device));
  for (int r = 0; r < nranks; ++r) {
    if (r == rank) {
      ctx->peer_mbox[r] = ctx->mailbox.p;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + (size_t)r * DOAPP_PEER_HANDLE_BYTES, sizeof(h));
    void *p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    ctx->peer_mbox[r] = (unsigned long long *)p;
  }
  ctx->peer_rank = rank;
  ctx->peer_nranks = nranks;
  ctx->peer_ready = true;
  ctx->peer_epoch = 0;
  return DOAPP_OK;
}

doapp_status doapp_comm_unique_id(uint8_t id[DOAPP_COMM_ID_BYTES]) {
  if (!id) return DOAPP_INVALID_ARG;
  std::string err;
  NcclApi *api = nccl_api(&err);
  if (!api) return fail(nullptr, DOAPP_NCCL_ERROR, "%s", err.c_str());
  static_assert(sizeof(ncclUniqueId) <= DOAPP_COMM_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId uid;
  ncclResult_t r = api->GetUniqueId(&uid);
  if (r != ncclSuccess) return fail(nullptr, DOAPP_NCCL_ERROR, "ncclGetUniqueId failed (%d)", (int)r);
  memset(id, 0, DOAPP_COMM_ID_BYTES);
  memcpy(id, &uid, sizeof(uid));
  return DOAPP_OK;
}

doapp_status doapp_comm_init(doapp_ctx *ctx, const uint8_t id[DOAPP_COMM_ID_BYTES], int32_t rank,
                             int32_t nranks) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!id || nranks < 1 || rank < 0 || rank >= nranks)
    return fail(ctx, DOAPP_INVALID_ARG, "doapp_comm_init: rank %d of %d", rank, nranks);
  if (ctx->comm) return fail(ctx, DOAPP_INVALID_ARG, "doapp_comm_init: already initialised");
  std::string err;
  NcclApi *api = nccl_api(&err);
  if (!api) return fail(ctx, DOAPP_NCCL_ERROR, "%s", err.c_str());
  CU(cudaSetDevice(ctx->cfg.device));
  ncclUniqueId uid;
  memcpy(&uid, id, sizeof(uid));
  ncclResult_t r = api->CommInitRank(&ctx->comm, nranks, uid, rank);
  if (r != ncclSuccess) {
    ctx->comm = nullptr;
    return fail(ctx, DOAPP_NCCL_ERROR, "ncclCommInitRank: %s",
                api->GetErrorString ? api->GetErrorString(r) : "error");
  }
  ctx->rank = rank;
  ctx->nranks = nranks;
  invalidate_graph(ctx);
  return DOAPP_OK;
}

doapp_status doapp_grid_broadcast(doapp_ctx *ctx, int32_t root) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!ctx->comm) return fail(ctx, DOAPP_NOT_READY, "doapp_grid_broadcast: no communicator");
  if (ctx->grid.kind == DOAPP_GRID_NONE) return fail(ctx, DOAPP_NOT_READY, "no grid allocated");
  NcclApi *api = nccl_api(nullptr);
  CU(cudaSetDevice(ctx->cfg.device));
  ncclResult_t r = api->Broadcast(ctx->grid_cells.p, ctx->grid_cells.p, ctx->grid_bytes, ncclUint8,
                                  root, ctx->comm, ctx->stream);
  if (r != ncclSuccess)
    return fail(ctx, DOAPP_NCCL_ERROR, "ncclBroadcast: %s",
                api->GetErrorString ? api->GetErrorString(r) : "error");
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

// --------------------------------------------------------------- introspection

uint64_t doapp_launch_count(const doapp_ctx *ctx) { return ctx ? ctx->launches : 0; }

doapp_status doapp_plan_stats(doapp_ctx *ctx, uint64_t *accepted_updates, uint64_t *cost_passes) {
  if (!ctx) return DOAPP_INVALID_ARG;
  CU(cudaSetDevice(ctx->cfg.device));
  unsigned long long h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  CU(cudaStreamSynchronize(ctx->stream));
  if (ctx->p2_ran) CU(cudaMemcpy(h, ctx->p2_stats.p, sizeof(h), cudaMemcpyDeviceToHost));
  if (accepted_updates) *accepted_updates = h[0];
  if (cost_passes) *cost_passes = h[1];
  if (getenv("DOAPP_P2_TRACE") && ctx->p2_ran) { // per-iteration stamps of the first CTA (SM clocks)
    std::vector<unsigned long long> st(256 * 8);
    CU(cudaMemcpy(st.data(), ctx->p2_stats.p + 8, st.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    double d[5] = {0, 0, 0, 0, 0};
    int cnt = 0;
    for (int it = 20; it < 90 && it < (int)ctx->plan_iters; ++it) {
      const unsigned long long *r = &st[(size_t)it * 8];
      if (!r[1] || !r[2] || !r[3] || !r[4] || !r[5]) continue;
      d[0] += (double)(r[1] - r[0]); // pass
      d[1] += (double)((long long)(r[2] - r[1])); // pass end -> fold sees tags
      d[2] += (double)(r[3] - r[2]); // fold
      d[3] += (double)((long long)(r[4] - r[3])); // publish -> all keys seen
      d[4] += (double)(r[5] - r[4]); // decide
      ++cnt;
    }
    if (cnt)
      fprintf(stderr, "p2 stamps (%d iterations): pass %.0f | end->tags seen %.0f | fold %.0f | publish->all keys %.0f | decide %.0f\n",
              cnt, d[0] / cnt, d[1] / cnt, d[2] / cnt, d[3] / cnt, d[4] / cnt);
  }
  if (getenv("DOAPP_P2_TRACE")) // SM clocks of the first CTA's warps, by phase
    fprintf(stderr, "p2 trace: passes %llu eval %llu wait %llu | fold %llu wait %llu | poll %llu decide %llu\n",
            h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
  return DOAPP_OK;
}
uint32_t doapp_plan_mode(const doapp_ctx *ctx) { return ctx ? ctx->last_plan_mode : 0; }

doapp_status doapp_debug_sincos(doapp_ctx *ctx, const float *angles, size_t count, float *sin_out,
                                float *cos_out) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (count == 0) return DOAPP_OK;
  if (!angles || !sin_out || !cos_out) return fail(ctx, DOAPP_INVALID_ARG, "null argument");
  CU(cudaSetDevice(ctx->cfg.device));
  ST(ensure(ctx, ctx->dbg_a, count));
  ST(ensure(ctx, ctx->dbg_b, count));
  ST(ensure(ctx, ctx->dbg_c, count));
  CU(cudaMemcpyAsync(ctx->dbg_a.p, angles, count * sizeof(float), cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(launch_debug_sincos(ctx->dbg_a.p, count, ctx->dbg_b.p, ctx->dbg_c.p, ctx->stream));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(sin_out, ctx->dbg_b.p, count * sizeof(float), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaMemcpyAsync(cos_out, ctx->dbg_c.p, count * sizeof(float), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

doapp_status doapp_debug_fk(doapp_ctx *ctx, const float *waypoints, size_t count,
                            float *points_out) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (count == 0) return DOAPP_OK;
  if (!waypoints || !points_out) return fail(ctx, DOAPP_INVALID_ARG, "null argument");
  CU(cudaSetDevice(ctx->cfg.device));
  const size_t d = ctx->cfg.waypoint_dim;
  ST(ensure(ctx, ctx->dbg_a, count * d));
  ST(ensure(ctx, ctx->dbg_b, count * 15));
  CU(cudaMemcpyAsync(ctx->dbg_a.p, waypoints, count * d * sizeof(float), cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(launch_debug_fk(ctx->dbg_a.p, count, (int)d, ctx->rpy_s, ctx->rpy_c, ctx->dbg_b.p,
                     ctx->stream));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(points_out, ctx->dbg_b.p, count * 15 * sizeof(float), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

doapp_status doapp_debug_fk_model(doapp_ctx *ctx, const float *waypoints, size_t count,
                                  float *centres_out) {
  if (!ctx) return DOAPP_INVALID_ARG;
  if (!ctx->fk_on) return fail(ctx, DOAPP_NOT_READY, "doapp_debug_fk_model: no FK model set");
  if (count == 0) return DOAPP_OK;
  if (!waypoints || !centres_out) return fail(ctx, DOAPP_INVALID_ARG, "null argument");
  CU(cudaSetDevice(ctx->cfg.device));
  const size_t d = ctx->cfg.waypoint_dim, per = (size_t)ctx->fk_spheres * 3;
  ST(ensure(ctx, ctx->dbg_a, count * d));
  ST(ensure(ctx, ctx->dbg_b, count * per));
  CU(cudaMemcpyAsync(ctx->dbg_a.p, waypoints, count * d * sizeof(float), cudaMemcpyHostToDevice,
                     ctx->stream));
  CU(launch_debug_fk_model(ctx->fk_dev.p, ctx->dbg_a.p, count, (int)d, ctx->fk_spheres,
                           ctx->dbg_b.p, ctx->stream));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(centres_out, ctx->dbg_b.p, count * per * sizeof(float),
                     cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DOAPP_OK;
}

} // extern "C"
