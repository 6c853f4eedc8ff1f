// ref_fk_harness.cu — pins stage 2 (forward kinematics + collision count) against the
// REFERENCE's own device code.  TEST INFRASTRUCTURE ONLY; contains no reference code.
//
// The helpers of generate_trajectories.cu:303-387 (init_rotx/y/z, init_rot_rpy, eye) are
// `static __device__`, so oracle/build_ref.sh compiles the reference source INTO this
// translation unit (`#include GT_SOURCE`) instead of linking it:
//   GT_SOURCE = oracle/_ref/gt_nohack.cu, which is /root/reference/perturbator/src/
//   generate_trajectories.cu with the single line :394 (`return 0; //Quick hack`) dropped by sed
//   (a build product under the git-ignored oracle/_ref/, never committed).
// Dumps (JSON, floats as uint32 bit patterns):
//   rot_angles / rot_x / rot_y / rot_z   init_rotx/y/z for a few thousand angles       (:303-352)
//   rot_rpy                               init_rot_rpy(0, -pi/2, 0)                     (:354-367,456)
//   coll_waypoints / coll_counts          num_collisions(waypoint, 5, grid) over random
//                                         waypoints on a planted distance grid         (:391-466)
//   coll_grid_dims / coll_grid_seed       the planted grid: cell (ix,iy,iz) holds 0.0 (collides)
//                                         or 1.0 by a hash of its logical index
// With the hack removed the chain still runs the as-shipped Matrix*Vector (row sums):
// that is what the reference's device code computes, and what orc_fk_points_as_shipped_matvec
// restates; the axes order (y,z,y,y,y), the rpy composition, the link order and the `<= 0.05f`
// predicate are all observable through it.
#define private public // test-only: reach CPUDistanceGrid's storage to plant the cells
#include "cpu_distance_grid.cuh"
#undef private
#include GT_SOURCE

#include <cmath>
#include <cstdio>
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

template <typename T> static T *managed(size_t n) {
  void *p = nullptr;
  if (cudaMallocManaged(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
  memset(p, 0, n * sizeof(T));
  return static_cast<T *>(p);
}

__global__ void k_rotations(const float *ang, int count, float *rx, float *ry, float *rz) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  Matrix<float, 3, 3> R(3, 3);
  init_rotx(R, ang[i]);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) rx[i * 9 + a * 3 + b] = R[a][b];
  init_roty(R, ang[i]);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) ry[i * 9 + a * 3 + b] = R[a][b];
  init_rotz(R, ang[i]);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) rz[i * 9 + a * 3 + b] = R[a][b];
}

__global__ void k_rpy(float *o) {
  Matrix<float, 3, 3> R(3, 3);
  init_rot_rpy(R, 0.0f, -1.5707963267948966f, 0.0f); // the call of :456
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) o[a * 3 + b] = R[a][b];
  eye(R);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) o[9 + a * 3 + b] = R[a][b];
}

__global__ void k_collisions(float *wps, int count, const doapp::CPUDistanceGrid *grid,
                             long long *counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  counts[i] = (long long)num_collisions(wps + (size_t)i * 5, 5, *grid);
}

static unsigned hash32(unsigned x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

int main(int argc, char **argv) {
  out = argc > 1 ? fopen(argv[1], "w") : stdout;
  if (!out) return 2;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    fprintf(stderr, "ref_fk_harness needs a GPU\n");
    return 3;
  }
  fputc('{', out);
  std::mt19937 rng(467);
  // ---- rotation builders
  const int A = 4096;
  float *ang = managed<float>(A), *rx = managed<float>(A * 9), *ry = managed<float>(A * 9),
        *rz = managed<float>(A * 9), *rpy = managed<float>(18);
  std::uniform_real_distribution<float> UA(-6.5f, 6.5f);
  for (int i = 0; i < A; ++i) ang[i] = UA(rng);
  ang[0] = 0.0f; ang[1] = 1.5707963267948966f; ang[2] = -1.5707963267948966f; ang[3] = 3.14159265f;
  k_rotations<<<(A + 127) / 128, 128>>>(ang, A, rx, ry, rz);
  k_rpy<<<1, 1>>>(rpy);
  if (cudaDeviceSynchronize() != cudaSuccess) return 4;
  put_f32("rot_angles", ang, A);
  put_f32("rot_x", rx, A * 9);
  put_f32("rot_y", ry, A * 9);
  put_f32("rot_z", rz, A * 9);
  put_f32("rot_rpy", rpy, 9);
  put_f32("rot_eye", rpy + 9, 9);
  // ---- num_collisions on a planted grid large enough for the as-shipped row sums
  // (|c + s| <= 1.42, z = c - s): 128^3 cells of 5 cm -> x,y in [-3.2, 3.2), z in [0, 6.4)
  const size_t N = 128;
  const double res = 0.05;
  const unsigned seed = 20191u;
  void *mem = nullptr;
  if (cudaMallocManaged(&mem, sizeof(doapp::CPUDistanceGrid)) != cudaSuccess) return 4;
  doapp::cpu_distance_grid::Dimensions dims{N, N, N, res};
  doapp::CPUDistanceGrid *g = new (mem) doapp::CPUDistanceGrid(dims);
  for (size_t z = 0; z < N; ++z)
    for (size_t y = 0; y < N; ++y)
      for (size_t x = 0; x < N; ++x) {
        const unsigned idx = (unsigned)(x + N * y + N * N * z);
        // 0.0 <= 0.05 collides, 1.0 does not; a third of the cells sit at the predicate's edge
        const unsigned h = hash32(idx ^ seed) % 6u;
        const float v = h == 0 ? 0.0f : h == 1 ? 0.05f : h == 2 ? 0.050000004f : 1.0f;
        g->aligned_base_[g->slice_pitch_ * z + g->dimensions_.length * y + x] = v;
      }
  const int W = 8192;
  float *wps = managed<float>((size_t)W * 5);
  long long *counts = managed<long long>(W);
  std::uniform_real_distribution<float> UW(-3.2f, 3.2f);
  for (int i = 0; i < W * 5; ++i) wps[i] = UW(rng);
  for (int j = 0; j < 5; ++j) wps[j] = 0.0f; // the zero pose
  k_collisions<<<(W + 127) / 128, 128>>>(wps, W, g, counts);
  if (cudaDeviceSynchronize() != cudaSuccess) {
    fprintf(stderr, "k_collisions: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 4;
  }
  put_f32("coll_waypoints", wps, (size_t)W * 5);
  put_i64("coll_counts", counts, W);
  long long gd[3] = {(long long)N, (long long)seed, 0};
  put_i64("coll_grid_n_seed", gd, 2);
  float gr[1] = {(float)res};
  put_f32("coll_grid_res", gr, 1);
  fprintf(out, "\n}\n");
  if (out != stdout) fclose(out);
  return 0;
}
