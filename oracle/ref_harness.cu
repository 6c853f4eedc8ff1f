// ref_harness.cu — runs the REFERENCE's own code (compiled from the sources where they lie
// under /root/reference by oracle/build_ref.sh) and dumps its outputs as JSON (float arrays
// as uint32 bit patterns).  TEST INFRASTRUCTURE ONLY.  This file contains no reference code:
// it only calls it.
//
//   part "host":   Matrix/Vector operators of perturbator/include/matrix.cuh on the host
//                  (the 2x2 golden product of perturbator/src/test_matrix.cu, 3x3 float
//                  products, the as-shipped Matrix*Vector)                    -> runs anywhere
//   part "grid":   CPUDistanceGrid ctor + operator() (cpu_distance_grid.cu:28-83,113-126)
//                  on a grid whose cells hold their own logical index        -> needs a GPU
//                  (the ctor calls cudaMallocManaged)
//   part "kernel": init_cudarand + one iteration of optimize_trajectories
//                  (generate_trajectories.cu:17-25,144-258) with the stop flag preset, as
//                  planning_state.cu:56-90 launches it                       -> needs a GPU
#define private public // test-only: reach CPUDistanceGrid's storage to plant cell ids
#include "cpu_distance_grid.cuh"
#undef private
#include "generate_trajectories.cuh"
#include "matrix.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <random>
#include <vector>

static FILE *out;
static bool first_key = true;
static void key(const char *k) {
  fprintf(out, "%s\n \"%s\": ", first_key ? "" : ",", k);
  first_key = false;
}
static void put_f32(const char *k, const float *p, size_t n) {
  key(k);
  fputc('[', out);
  for (size_t i = 0; i < n; ++i) {
    unsigned u;
    memcpy(&u, p + i, 4);
    fprintf(out, "%s%u", i ? "," : "", u);
  }
  fputc(']', out);
}
static void put_i64(const char *k, const long long *p, size_t n) {
  key(k);
  fputc('[', out);
  for (size_t i = 0; i < n; ++i) fprintf(out, "%s%lld", i ? "," : "", p[i]);
  fputc(']', out);
}

static void part_host() {
  using doapp::Matrix;
  using doapp::Vector;
  Matrix<int, 2, 2> A(2, 2), B(2, 2);
  A[0][0] = 1; A[0][1] = 2; A[1][0] = 3; A[1][1] = 4;
  B[0][0] = 5; B[0][1] = 6; B[1][0] = 7; B[1][1] = 8;
  const Matrix<int, 2, 2> C = A * B;
  long long c[4] = {C[0][0], C[0][1], C[1][0], C[1][1]};
  put_i64("matmul_2x2_int", c, 4);

  std::mt19937 rng(467);
  std::uniform_real_distribution<float> U(-2.0f, 2.0f);
  const int T = 64;
  std::vector<float> a(T * 9), b(T * 9), v(T * 3), ab(T * 9), av(T * 3);
  for (int t = 0; t < T; ++t) {
    Matrix<float, 3, 3> Ma(3, 3), Mb(3, 3);
    Vector<float, 3> Vv(3);
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) {
        Ma[i][j] = a[t * 9 + i * 3 + j] = U(rng);
        Mb[i][j] = b[t * 9 + i * 3 + j] = U(rng);
      }
      Vv[i] = v[t * 3 + i] = U(rng);
    }
    const Matrix<float, 3, 3> Mc = Ma * Mb;
    const Vector<float, 3> Vu = Ma * Vv; // as shipped: ignores v
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) ab[t * 9 + i * 3 + j] = Mc[i][j];
      av[t * 3 + i] = Vu[i];
    }
  }
  put_f32("mat_a", a.data(), a.size());
  put_f32("mat_b", b.data(), b.size());
  put_f32("vec_v", v.data(), v.size());
  put_f32("mat_ab", ab.data(), ab.size());
  put_f32("mat_av_as_shipped", av.data(), av.size());
}

template <typename T> static T *managed(size_t n) {
  void *p = nullptr;
  if (cudaMallocManaged(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
  memset(p, 0, n * sizeof(T));
  return static_cast<T *>(p);
}

static doapp::CPUDistanceGrid *make_grid(size_t L, size_t W, size_t H, double res) {
  void *mem = nullptr; // the kernel takes `const CPUDistanceGrid &`: the object itself must be
  if (cudaMallocManaged(&mem, sizeof(doapp::CPUDistanceGrid)) != cudaSuccess) return nullptr;
  doapp::cpu_distance_grid::Dimensions dims{L, W, H, res}; // device-visible (unique_ptr.cuh:76-93)
  try {
    return new (mem) doapp::CPUDistanceGrid(dims);
  } catch (...) {
    return nullptr;
  }
}

static bool part_grid(const char *tag, size_t N, double res) {
  doapp::CPUDistanceGrid *g = make_grid(N, N, N, res);
  if (!g) return false;
  for (size_t z = 0; z < N; ++z)
    for (size_t y = 0; y < N; ++y)
      for (size_t x = 0; x < N; ++x) // cell value = its logical index ix + N*iy + N*N*iz
        g->aligned_base_[g->slice_pitch_ * z + g->dimensions_.length * y + x] =
            float(x + N * y + N * N * z);
  std::mt19937 rng(2019);
  const float ext = float(N * res);
  std::uniform_real_distribution<float> U(-0.6f * ext, 1.1f * ext);
  const int P = 6000;
  std::vector<float> pts, vals;
  for (int i = 0; i < P; ++i) {
    float x = U(rng), y = U(rng), z = U(rng);
    if (i % 3 == 0) { // exact multiples of the resolution: cell boundaries
      x = float(res) * float(int(x / float(res)));
      y = float(res) * float(int(y / float(res)));
      z = float(res) * float(int(z / float(res)));
    }
    pts.push_back(x); pts.push_back(y); pts.push_back(z);
    // coord == max passes the reference's inclusive bounds test and indexes one past the
    // end: undefined behaviour, not called; recorded as NaN
    const bool ub = x == g->x_max_ || y == g->y_max_ || z == g->z_max_;
    vals.push_back(ub ? NAN : (*g)(x, y, z));
  }
  char name[64];
  snprintf(name, sizeof name, "grid_%s_points", tag);
  put_f32(name, pts.data(), pts.size());
  snprintf(name, sizeof name, "grid_%s_values", tag);
  put_f32(name, vals.data(), vals.size());
  float frame[5] = {g->resolution_, g->x_offset_, g->y_offset_, g->z_max_, float(g->slice_pitch_)};
  snprintf(name, sizeof name, "grid_%s_frame", tag);
  put_f32(name, frame, 5);
  return true;
}

static bool part_kernel() {
  const unsigned k = 10, m = 5, n = 50, d = doapp::num_joints; // n*m*d = 1000 threads (D18)
  const float dt = 0.1f;                                       // planning_state.cuh:55
  const unsigned per_traj = d * (n > m ? n : m), num_rngs = k * per_traj;
  doapp::CPUDistanceGrid *g = make_grid(100, 100, 100, 0.01); // main.cu:18
  if (!g) return false;
  float *traj = managed<float>(k * n * d), *noise = managed<float>(k * m * d);
  float *noisy = managed<float>(k * m * n * d), *vel = managed<float>(k * m * n * d);
  float *acc = managed<float>(k * m * n * d), *smooth = managed<float>(k * m);
  float *scores = managed<float>(k * (m + 1)), *upd = managed<float>(k * d);
  bool *found = managed<bool>(k), *flag = managed<bool>(1);
  curandState *rngs = managed<curandState>(num_rngs);
  if (!traj || !rngs || !flag) return false;
  init_cudarand<<<(num_rngs + 511) / 512, 512>>>(rngs, num_rngs); // planning_state.cu:33-36
  if (cudaDeviceSynchronize() != cudaSuccess) return false;
  const float start[4] = {0.1f, -0.2f, 0.3f, 0.05f}, goal[4] = {0.0f, 1.0f, 0.0f, 0.5f};
  std::vector<float> lerp(n * d);
  for (unsigned w = 0; w < n; ++w)
    for (unsigned j = 0; j < d; ++j)
      lerp[w * d + j] = start[j] + float(w + 1) * ((goal[j] - start[j]) / float(n));
  cudaMemcpyToSymbol(initial_waypoint, start, sizeof start);
  cudaMemcpyToSymbol(final_waypoint, goal, sizeof goal);
  cudaMemcpyToSymbol(initial_trajectory, lerp.data(), lerp.size() * sizeof(float));
  flag[0] = true; // do { ... } while(!*time_expired): exactly one iteration
  optimize_trajectories<<<k, n * m * d>>>(traj, noise, noisy, rngs, vel, acc, smooth, scores, upd,
                                          found, flag, per_traj, n, d, m, dt, *g);
  if (cudaDeviceSynchronize() != cudaSuccess) {
    fprintf(stderr, "optimize_trajectories: %s\n", cudaGetErrorString(cudaGetLastError()));
    return false;
  }
  long long shape[4] = {k, m, n, d};
  put_i64("kernel_shape_kmnd", shape, 4);
  put_f32("kernel_start", start, 4);
  put_f32("kernel_initial_trajectory", lerp.data(), lerp.size());
  put_f32("kernel_noise", noise, k * m * d);
  put_f32("kernel_noisy", noisy, k * m * n * d);
  put_f32("kernel_vel", vel, k * m * n * d);
  put_f32("kernel_acc", acc, k * m * n * d);
  put_f32("kernel_smoothness", smooth, k * m);
  put_f32("kernel_scores", scores, k * (m + 1));
  return true;
}

// --time N: the reference kernel at the shape the shipped node launches (main.cu:20-21,
// planning_state.cu:61-62,79-84: k = 1000 blocks x n*m*d = 1000 threads), one optimiser iteration
// per launch (stop flag preset), N launches timed with CUDA events after 3 warm-ups.  The
// kernel's own device printf (gt.cu:194,237,255) is part of what is timed: it is the reference
// as shipped.  One JSON line on stdout (after the kernel's prints).
static int part_time(int reps) {
  const unsigned k = 1000, n = 50, d = doapp::num_joints, m = 1024 / n / d; // main.cu:20-21
  const float dt = 0.1f;
  const unsigned per_traj = d * (n > m ? n : m), num_rngs = k * per_traj;
  doapp::CPUDistanceGrid *g = make_grid(100, 100, 100, 0.01);
  if (!g) return 3;
  float *traj = managed<float>(k * n * d), *noise = managed<float>(k * m * d);
  float *noisy = managed<float>(k * m * n * d), *vel = managed<float>(k * m * n * d);
  float *acc = managed<float>(k * m * n * d), *smooth = managed<float>(k * m);
  float *scores = managed<float>(k * (m + 1)), *upd = managed<float>(k * d);
  bool *found = managed<bool>(k), *flag = managed<bool>(1);
  curandState *rngs = managed<curandState>(num_rngs);
  if (!traj || !rngs || !flag) return 3;
  init_cudarand<<<(num_rngs + 511) / 512, 512>>>(rngs, num_rngs);
  if (cudaDeviceSynchronize() != cudaSuccess) return 3;
  const float start[4] = {0.0f, 0.0f, 0.0f, 0.0f}, goal[4] = {0.0f, -1.0f, 1.5f, -1.5f};
  std::vector<float> lerp(n * d);
  for (unsigned w = 0; w < n; ++w)
    for (unsigned j = 0; j < d; ++j)
      lerp[w * d + j] = start[j] + float(w + 1) * ((goal[j] - start[j]) / float(n));
  cudaMemcpyToSymbol(initial_waypoint, start, sizeof start);
  cudaMemcpyToSymbol(final_waypoint, goal, sizeof goal);
  cudaMemcpyToSymbol(initial_trajectory, lerp.data(), lerp.size() * sizeof(float));
  flag[0] = true;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float ms = 0.0f;
  for (int r = -3; r < reps; ++r) {
    if (r == 0) cudaEventRecord(e0);
    optimize_trajectories<<<k, n * m * d>>>(traj, noise, noisy, rngs, vel, acc, smooth, scores, upd,
                                            found, flag, per_traj, n, d, m, dt, *g);
  }
  cudaEventRecord(e1);
  if (cudaDeviceSynchronize() != cudaSuccess) {
    fprintf(stderr, "optimize_trajectories: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 3;
  }
  cudaEventElapsedTime(&ms, e0, e1);
  const double wps = double(k) * m * n, us = 1e3 * ms / reps;
  fflush(stdout);
  printf("{\"kernel\": \"optimize_trajectories (reference, as shipped: collision term stubbed)\", "
         "\"shape_kmnd\": [%u, %u, %u, %u], \"launches\": %d, \"us_per_iteration\": %.3f, "
         "\"waypoint_samples_per_iteration\": %.0f, \"value\": %.6g, "
         "\"unit\": \"waypoint-samples/s\"}\n",
         k, m, n, d, reps, us, wps, wps / (us * 1e-6));
  return 0;
}

int main(int argc, char **argv) {
  if (argc > 2 && !strcmp(argv[1], "--time")) return part_time(atoi(argv[2]) > 0 ? atoi(argv[2]) : 20);
  out = argc > 1 ? fopen(argv[1], "w") : stdout;
  if (!out) return 2;
  fputc('{', out);
  part_host();
  int ndev = 0;
  bool gpu = cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0;
  long long have[3] = {1, 0, 0};
  if (gpu) {
    have[1] = part_grid("100", 100, 0.01) && part_grid("64", 64, 0.02) && part_grid("128", 128, 0.01);
    have[2] = part_kernel();
  }
  put_i64("parts_host_grid_kernel", have, 3);
  fprintf(out, "\n}\n");
  if (out != stdout) fclose(out);
  return 0;
}
