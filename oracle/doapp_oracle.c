/*
 * doapp_oracle.c — CPU restatement of the reference's perturbator cost path.
 * TEST INFRASTRUCTURE ONLY (see doapp_oracle.h).  Build: oracle/Makefile
 * (gcc -O3 -march=x86-64-v3 -ffp-contract=off: every rounding is written
 * explicitly; fmaf() below is a single-rounding fused multiply-add).
 */
#include "doapp_oracle.h"

#include <limits.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

void orc_params_init(orc_params *p) {
  memset(p, 0, sizeof(*p));
  p->k = 1;
  p->m = 1024;
  p->n = 64;
  p->d = 5;
  p->n_noise = 1;
  p->delta_t = 0.1f;            /* planning_state.cuh:55 */
  p->smooth_scale = 0.25f;      /* gt.cu:81,107: two `*= 0.5f` */
  p->collision_distance = 0.05f; /* gt.cu:389 */
  p->collision_mode = ORC_COLLISION_INTENDED;
  p->select_mode = ORC_SELECT_CEIL_INT;
  p->sum_order = ORC_SUM_WAYPOINT;
  p->sigma = 1.0f;
  p->seed = 1234; /* gt.cu:23 */
  p->update_mode = ORC_UPDATE_KEEP_BEST;
  p->update_gain = 10000.0f; /* gt.cu:140 */
}

/* ------------------------------------------------------------------------
 * Deterministic elementary functions.  The reference calls std::sin/std::cos
 * on the device (gt.cu:304-305); CUDA's and glibc's sinf differ in the last
 * ulp, which would flip voxel indices, so both sides of the parity check use
 * this specification instead (DESIGN.md "deterministic math spec"):
 *   |x| < 2^17:  t = x*(2/pi) [RN]; kf = rintf(t);
 *                r = fma(-kf,P1,x); r = fma(-kf,P2,r); r = fma(-kf,P3,r)
 *   otherwise :  xd = fmod(x, 2*pi) in binary64 (exact), then the two-constant
 *                reduction in binary64, r -> float
 *   z = r*r
 *   sin_r = fma(fma(fma(S3,z,S2),z,S1)*z, r, r)
 *   cos_r = fma(fma(fma(fma(C3,z,C2),z,C1),z,-0.5), z, 1)
 *   quadrant q = k & 3 : (s,c) = (sin_r,cos_r), (cos_r,-sin_r), (-sin_r,-cos_r), (-cos_r,sin_r)
 * Polynomial coefficients: Cephes sinf/cosf minimax on [-pi/4, pi/4].
 * ---------------------------------------------------------------------- */
static const float ORC_TWO_OVER_PI = 0x1.45f306p-1f;
static const float ORC_P1 = 0x1.921fb6p+0f;    /* RN(pi/2)               */
static const float ORC_P2 = -0x1.777a5cp-25f;  /* RN(pi/2 - P1)          */
static const float ORC_P3 = -0x1.ee59dap-50f;  /* RN(pi/2 - P1 - P2)     */
static const double ORC_TWO_OVER_PI_D = 0x1.45f306dc9c883p-1;
static const double ORC_TWO_PI_D = 0x1.921fb54442d18p+2;
static const double ORC_P1_D = 0x1.921fb54442d18p+0;
static const double ORC_P2_D = 0x1.1a62633145c07p-54;
static const float ORC_S1 = -1.6666654611e-1f;
static const float ORC_S2 = 8.3321608736e-3f;
static const float ORC_S3 = -1.9515295891e-4f;
static const float ORC_C1 = 4.166664568298827e-2f;
static const float ORC_C2 = -1.388731625493765e-3f;
static const float ORC_C3 = 2.443315711809948e-5f;

void orc_sincosf(float x, float *s, float *c) {
  float r;
  int q;
  if (fabsf(x) < 131072.0f) {
    float t = x * ORC_TWO_OVER_PI;
    float kf = rintf(t);
    r = fmaf(-kf, ORC_P1, x);
    r = fmaf(-kf, ORC_P2, r);
    r = fmaf(-kf, ORC_P3, r);
    q = (int)kf & 3;
  } else if (x != x || fabsf(x) == INFINITY) {
    *s = NAN;
    *c = NAN;
    return;
  } else {
    /* exact remainder by the binary64 value of 2*pi keeps the argument small;
     * accuracy decays with |x| but the result stays bounded and deterministic */
    double xd = fmod((double)x, ORC_TWO_PI_D);
    double kd = rint(xd * ORC_TWO_OVER_PI_D);
    double rd = fma(-kd, ORC_P1_D, xd);
    rd = fma(-kd, ORC_P2_D, rd);
    r = (float)rd;
    q = (int)kd & 3;
  }
  float z = r * r;
  float ps = fmaf(ORC_S3, z, ORC_S2);
  ps = fmaf(ps, z, ORC_S1);
  ps = ps * z;
  float sin_r = fmaf(ps, r, r);
  float pc = fmaf(ORC_C3, z, ORC_C2);
  pc = fmaf(pc, z, ORC_C1);
  pc = fmaf(pc, z, -0.5f);
  float cos_r = fmaf(pc, z, 1.0f);
  switch (q) {
  case 0: *s = sin_r; *c = cos_r; break;
  case 1: *s = cos_r; *c = -sin_r; break;
  case 2: *s = -sin_r; *c = -cos_r; break;
  default: *s = -cos_r; *c = sin_r; break;
  }
}

/* natural log for x in (0, 1] (Box-Muller radius), Cephes logf polynomial on
 * [sqrt(1/2), sqrt(2)) evaluated with explicit fmaf. */
float orc_logf(float x) {
  uint32_t ix;
  memcpy(&ix, &x, 4);
  ix += 0x3f800000u - 0x3f3504f3u;
  int e = (int)(ix >> 23) - 127;
  ix = (ix & 0x007fffffu) + 0x3f3504f3u;
  float mant;
  memcpy(&mant, &ix, 4);
  float f = mant - 1.0f;
  float z = f * f;
  float p = 7.0376836292e-2f;
  p = fmaf(p, f, -1.1514610310e-1f);
  p = fmaf(p, f, 1.1676998740e-1f);
  p = fmaf(p, f, -1.2420140846e-1f);
  p = fmaf(p, f, 1.4249322787e-1f);
  p = fmaf(p, f, -1.6668057665e-1f);
  p = fmaf(p, f, 2.0000714765e-1f);
  p = fmaf(p, f, -2.4999993993e-1f);
  p = fmaf(p, f, 3.3333331174e-1f);
  float y = (f * z) * p;
  float fe = (float)e;
  y = fmaf(fe, -2.12194440e-4f, y);
  y = fmaf(-0.5f, z, y);
  float r = f + y;
  r = fmaf(fe, 0.693359375f, r);
  return r;
}

/* ------------------------------------------------------------------------
 * Stage 1.  Replaces init_cudarand / curand_normal (gt.cu:17-25,37-44,297-299;
 * XORWOW, not reproduced: the reference's tests pin nothing about the stream).
 * Philox4x32-10, key = (seed lo, seed hi), counter =
 * (t_global, climber, noise waypoint, 2*iteration + call); four outputs give
 * two Box-Muller pairs; joint j uses call j/4, pair (j%4)/2, cos for even j.
 * ---------------------------------------------------------------------- */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static float orc_box_muller(uint32_t a, uint32_t b, int want_sin, float sigma) {
  float u1 = (float)((a >> 8) + 1u) * 0x1p-24f; /* (0,1] */
  float u2 = (float)(b >> 8) * 0x1p-24f;        /* [0,1) */
  float rad = sqrtf(-2.0f * orc_logf(u1));
  float s, c;
  orc_sincosf(u2 * 0x1.921fb6p+2f, &s, &c);
  float zn = rad * (want_sin ? s : c);
  return sigma * zn;
}

float orc_noise_at(const orc_params *p, uint32_t iteration, uint32_t c, uint64_t t_global,
                   uint32_t wn, uint32_t j) {
  uint32_t ctr[4] = {(uint32_t)t_global, c, wn, 2u * iteration + (j >> 2)};
  uint32_t key[2] = {(uint32_t)p->seed, (uint32_t)(p->seed >> 32)};
  uint32_t r[4];
  orc_philox4x32_10(ctr, key, r);
  uint32_t pair = (j & 3u) >> 1;
  return orc_box_muller(r[2 * pair], r[2 * pair + 1], (int)(j & 1u), p->sigma);
}

void orc_noise_fill(const orc_params *p, uint32_t iteration, float *noise) {
  for (uint32_t c = 0; c < p->k; ++c)
    for (uint32_t t = 0; t < p->m; ++t)
      for (uint32_t wn = 0; wn < p->n_noise; ++wn)
        for (uint32_t j = 0; j < p->d; ++j)
          noise[(((size_t)c * p->m + t) * p->n_noise + wn) * p->d + j] =
              orc_noise_at(p, iteration, c, p->sample_offset + t, wn, j);
}

/* ------------------------------------------------------------------------
 * Stage 2: forward kinematics, gt.cu:303-387 (rotation builders) and
 * gt.cu:397-463 (chain).  Matrix products accumulate like matrix.cuh:274-313:
 * c = T(); c += a[k]*b[k] for k = 0..2, which nvcc contracts to one FMA per
 * term; the restatement writes those FMAs explicitly.  The mat-vec product is
 * the INTENDED u[i] += a[i][j]*v[j] (matrix.cuh:315-336 ignores v).
 * ---------------------------------------------------------------------- */
typedef struct { float a[3][3]; } orc_m3;
typedef struct { float v[3]; } orc_v3;

static orc_m3 orc_rotx(float th) { /* gt.cu:303-318 */
  float s, c;
  orc_sincosf(th, &s, &c);
  orc_m3 R = {{{1.0f, 0.0f, 0.0f}, {0.0f, c, -s}, {0.0f, s, c}}};
  return R;
}
static orc_m3 orc_roty(float th) { /* gt.cu:320-335 */
  float s, c;
  orc_sincosf(th, &s, &c);
  orc_m3 R = {{{c, 0.0f, s}, {0.0f, 1.0f, 0.0f}, {-s, 0.0f, c}}};
  return R;
}
static orc_m3 orc_rotz(float th) { /* gt.cu:337-352 */
  float s, c;
  orc_sincosf(th, &s, &c);
  orc_m3 R = {{{c, -s, 0.0f}, {s, c, 0.0f}, {0.0f, 0.0f, 1.0f}}};
  return R;
}
static orc_m3 orc_eye(void) { /* gt.cu:375-387 */
  orc_m3 R = {{{1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}}};
  return R;
}
static orc_m3 orc_matmul(const orc_m3 *A, const orc_m3 *B) { /* matrix.cuh:274-313 */
  orc_m3 C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      float acc = 0.0f;
      for (int k = 0; k < 3; ++k) acc = fmaf(A->a[i][k], B->a[k][j], acc);
      C.a[i][j] = acc;
    }
  return C;
}
static orc_v3 orc_matvec(const orc_m3 *A, const orc_v3 *v) { /* matrix.cuh:315-336, D2 fixed */
  orc_v3 u;
  for (int i = 0; i < 3; ++i) {
    float acc = 0.0f;
    for (int j = 0; j < 3; ++j) acc = fmaf(A->a[i][j], v->v[j], acc);
    u.v[i] = acc;
  }
  return u;
}
static orc_v3 orc_matvec_as_shipped(const orc_m3 *A, const orc_v3 *v) { /* matrix.cuh:315-336 */
  (void)v;
  orc_v3 u;
  for (int i = 0; i < 3; ++i) {
    float acc = 0.0f;
    for (int j = 0; j < 3; ++j) acc += A->a[i][j];
    u.v[i] = acc;
  }
  return u;
}
static orc_v3 orc_vadd(const orc_v3 *a, const orc_v3 *b) { /* vector.cuh:253-264 */
  orc_v3 w = {{a->v[0] + b->v[0], a->v[1] + b->v[1], a->v[2] + b->v[2]}};
  return w;
}
static orc_m3 orc_rot_rpy(float roll, float pitch, float yaw) { /* gt.cu:354-367 */
  orc_m3 R = orc_rotx(roll);
  orc_m3 cur = orc_roty(pitch);
  R = orc_matmul(&R, &cur);
  cur = orc_rotz(yaw);
  R = orc_matmul(&R, &cur);
  return R;
}

static void orc_fk_impl(const float *waypoint, uint32_t d, float out[5][3], int shipped_matvec) {
  /* link translations, gt.cu:407,418,429,440,451 (= rexarm.urdf joint origins) */
  static const float T[5][3] = {{0.0f, -0.0015f, 0.0715f},
                                {0.0f, 0.0f, 0.045f},
                                {0.0f, 0.0f, 0.1f},
                                {0.0f, 0.0f, 0.1f},
                                {0.0f, -0.0125f, 0.08f}};
  orc_m3 stat = orc_eye();          /* gt.cu:406 */
  orc_v3 pos = {{0.0f, 0.0f, 0.0f}}; /* Vector(3) value-initialises, vector.cuh:79-88 */
  for (int link = 0; link < 5; ++link) {
    /* a 4-joint waypoint has no wrist_rotate angle (the reference reads
     * waypoint[4] out of bounds): use 0. */
    float th = (uint32_t)link < d ? waypoint[link] : 0.0f;
    orc_m3 dyn = (link == 1) ? orc_rotz(th) : orc_roty(th); /* axes y,z,y,y,y */
    orc_v3 t = {{T[link][0], T[link][1], T[link][2]}};
    orc_v3 sp = shipped_matvec ? orc_matvec_as_shipped(&stat, &pos) : orc_matvec(&stat, &pos);
    orc_v3 q = orc_vadd(&sp, &t);
    pos = shipped_matvec ? orc_matvec_as_shipped(&dyn, &q) : orc_matvec(&dyn, &q);
    if (link == 4) {
      /* gt.cu:456-459 */
      stat = orc_rot_rpy(0.0f, -1.5707963267948966f, 0.0f);
      orc_v3 t2 = {{0.0f, -0.005f, 0.07f}};
      orc_v3 sp2 = shipped_matvec ? orc_matvec_as_shipped(&stat, &pos) : orc_matvec(&stat, &pos);
      pos = orc_vadd(&sp2, &t2);
    }
    out[link][0] = pos.v[0];
    out[link][1] = pos.v[1];
    out[link][2] = pos.v[2];
  }
}

/* test access to the rotation builders (pinned against the reference's own device code,
 * tests/golden/ref_fk_outputs.npz): axis 0/1/2 = gt.cu:303-318 / 320-335 / 337-352 */
void orc_rotation(int axis, float theta, float out[9]) {
  orc_m3 R = axis == 0 ? orc_rotx(theta) : axis == 1 ? orc_roty(theta) : orc_rotz(theta);
  memcpy(out, R.a, sizeof(R.a));
}
void orc_rotation_rpy(float roll, float pitch, float yaw, float out[9]) { /* gt.cu:354-367 */
  orc_m3 R = orc_rot_rpy(roll, pitch, yaw);
  memcpy(out, R.a, sizeof(R.a));
}

void orc_fk_points(const float *waypoint, uint32_t d, float out[5][3]) {
  orc_fk_impl(waypoint, d, out, 0);
}
void orc_fk_points_as_shipped_matvec(const float *waypoint, uint32_t d, float out[5][3]) {
  orc_fk_impl(waypoint, d, out, 1);
}

/* ------------------------------------------------------------------------
 * URDF-correct FK with per-link spheres.  Not in the
 * reference's code: its chain (above) rotates a point, base->tip, about the
 * wrong axes (D14).  Here the link frames accumulate as the URDF defines them,
 *   p_i = p_{i-1} + R_{i-1} * origin_i,   R_i = R_{i-1} * Rot(axis_i, theta_i),
 * and a sphere of link i sits at p_i + R_i * centre.  Arithmetic spec (shared
 * with the CUDA kernels): R is kept as three columns; R * v is
 * fma(col2, v.z, fma(col1, v.y, col0 * v.x)) per component; a rotation about
 * axis a with cyclic successors (b, c) replaces
 *   col_b <- fma(s, col_c, c * col_b),  col_c <- fma(c, col_c, (-s) * col_b).
 * ---------------------------------------------------------------------- */
void orc_fk_model_rexarm(orc_fk_model *m) {
  static const float origin[5][3] = {{0.0f, -0.0015f, 0.0715f}, /* rexarm.urdf:64 */
                                     {0.0f, 0.0f, 0.045f},      /* :83 */
                                     {0.0f, 0.0f, 0.1f},        /* :109 */
                                     {0.0f, 0.0f, 0.1f},        /* :128 */
                                     {0.0f, -0.0125f, 0.08f}};  /* :147 */
  static const int32_t axis[5] = {2, 1, 1, 1, 2};               /* :65,84,110,129,148 */
  /* two spheres along the long side of every link box (centre -+ a quarter of the long
   * side, radius = half the larger short side), two along the fingers, one at the
   * gripper centre */
  static const orc_fk_sphere sph[13] = {
      {0, {0.0f, 0.0f, 0.0185f}, 0.0175f},      {0, {0.0f, 0.0f, 0.0435f}, 0.0175f},     /* :52,55 */
      {1, {0.0f, 0.0f, 0.0735f}, 0.0175f},      {1, {0.0f, 0.0f, 0.0985f}, 0.0175f},     /* :71,74 */
      {2, {0.0f, 0.0f, 0.07325f}, 0.016f},      {2, {0.0f, 0.0f, 0.09825f}, 0.016f},     /* :97,100 */
      {3, {0.0f, -0.01425f, 0.0645f}, 0.016f},  {3, {0.0f, 0.01075f, 0.0645f}, 0.016f},  /* :116,119 */
      {4, {0.0f, -0.01125f, 0.019f}, 0.0155f},  {4, {0.0f, 0.01375f, 0.019f}, 0.0155f},  /* :135,138 */
      {4, {0.0f, 0.01f, 0.0525f}, 0.01375f},    {4, {0.0f, 0.01f, 0.0875f}, 0.01375f},   /* :180,189 */
      {4, {0.0f, -0.005f, 0.07f}, 0.01f}};                                               /* :197 */
  memset(m, 0, sizeof(*m));
  memcpy(m->joint_origin, origin, sizeof(origin));
  memcpy(m->joint_axis, axis, sizeof(axis));
  m->num_spheres = 13;
  memcpy(m->spheres, sph, sizeof(sph));
}

static void orc_frame_apply(float col[3][3], const float pos[3], const float v[3], float out[3]) {
  for (int i = 0; i < 3; ++i) {
    float t = col[0][i] * v[0];
    t = fmaf(col[1][i], v[1], t);
    t = fmaf(col[2][i], v[2], t);
    out[i] = pos[i] + t;
  }
}

void orc_fk_model_points(const orc_fk_model *m, const float *waypoint, uint32_t d, float (*out)[3]) {
  float col[3][3] = {{1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}}; /* col[j][i] = R[i][j] */
  float pos[3] = {0.0f, 0.0f, 0.0f};
  for (int link = 0; link < 5; ++link) {
    float np[3];
    orc_frame_apply(col, pos, m->joint_origin[link], np);
    memcpy(pos, np, sizeof(pos));
    float th = (uint32_t)link < d ? waypoint[link] : 0.0f, s, c;
    orc_sincosf(th, &s, &c);
    int a = m->joint_axis[link], b = (a + 1) % 3, cc = (a + 2) % 3;
    for (int i = 0; i < 3; ++i) {
      float cb = col[b][i], cv = col[cc][i];
      col[b][i] = fmaf(s, cv, c * cb);
      col[cc][i] = fmaf(c, cv, (-s) * cb);
    }
    for (uint32_t q = 0; q < m->num_spheres; ++q)
      if (m->spheres[q].link == link) orc_frame_apply(col, pos, m->spheres[q].centre, out[q]);
  }
}

/* ------------------------------------------------------------------------
 * Stage 3: grid lookup.  CDG frame: cdg.cu:69-82 (offsets) and :113-126
 * (bounds, fp32 add, IEEE fp32 divide, truncation).  OG frame: og.cpp:29-40
 * (fp64).  Deviations: dense pitch; half-open bounds, and an
 * index that still lands on n (rounding of x + offset) is outside.
 * ---------------------------------------------------------------------- */
int64_t orc_grid_index(const orc_grid *g, float x, float y, float z) {
  if (g->frame == ORC_FRAME_CDG) {
    float res = (float)g->resolution;                            /* cdg.cu:69 */
    float x_off = (float)(0.5 * (double)g->nx * g->resolution);  /* cdg.cu:70-71 */
    float y_off = (float)(0.5 * (double)g->ny * g->resolution);  /* cdg.cu:72-73 */
    float z_max = (float)((double)g->nz * g->resolution);        /* cdg.cu:81-82 */
    if (!(x >= -x_off && x < x_off) || !(y >= -y_off && y < y_off) || !(z >= 0.0f && z < z_max))
      return -1;
    float fx = (x + x_off) / res; /* cdg.cu:120 */
    float fy = (y + y_off) / res; /* cdg.cu:121 */
    float fz = z / res;           /* cdg.cu:122 */
    uint64_t ix = (uint64_t)fx, iy = (uint64_t)fy, iz = (uint64_t)fz;
    if (ix >= g->nx || iy >= g->ny || iz >= g->nz) return -1;
    return (int64_t)((uint64_t)g->nx * g->ny * iz + (uint64_t)g->nx * iy + ix); /* cdg.cu:124-125 */
  } else {
    double size = (double)g->nx * g->resolution;
    double half = size / 2.0; /* og.cpp:31-33 */
    double xs = (double)x + half, ys = (double)y + half, zs = (double)z + half;
    if (!(xs >= 0.0 && ys >= 0.0 && zs >= 0.0)) return -1;
    double fx = xs / g->resolution, fy = ys / g->resolution, fz = zs / g->resolution; /* og.cpp:35-37 */
    if (!(fx < (double)g->nx && fy < (double)g->ny && fz < (double)g->nz)) return -1;
    uint64_t ix = (uint64_t)fx, iy = (uint64_t)fy, iz = (uint64_t)fz;
    if (ix >= g->nx || iy >= g->ny || iz >= g->nz) return -1;
    return (int64_t)(ix + (uint64_t)g->nx * iy + (uint64_t)g->nx * g->ny * iz); /* og.cpp:26 */
  }
}

int orc_grid_collides(const orc_grid *g, float x, float y, float z, float collision_distance) {
  if (g->kind == ORC_GRID_NONE || !g->cells) return 0;
  int64_t idx = orc_grid_index(g, x, y, z);
  if (idx < 0) return 0; /* HUGE_VALF <= 0.05f is false, cdg.cu:117 */
  if (g->kind == ORC_GRID_U8) return ((const uint8_t *)g->cells)[idx] != 0;
  return ((const float *)g->cells)[idx] <= collision_distance; /* gt.cu:412 */
}

size_t orc_grid_scatter(orc_grid *g, const float *xyz, size_t num_points, int8_t val,
                        int clear_first) {
  size_t cells = (size_t)g->nx * g->ny * g->nz;
  if (clear_first) memset(g->cells, 0, cells); /* og.cpp:15-16 */
  size_t inside = 0;
  for (size_t i = 0; i < num_points; ++i) {
    int64_t idx = orc_grid_index(g, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]); /* og.cpp:47-57 */
    if (idx < 0) continue;
    ((int8_t *)g->cells)[idx] = val; /* og.cpp:44 */
    ++inside;
  }
  return inside;
}

uint32_t orc_num_collisions(const orc_params *p, const orc_grid *g, const float *waypoint) {
  if (p->collision_mode == ORC_COLLISION_AS_SHIPPED) return 0; /* gt.cu:394 */
  if (!g || g->kind == ORC_GRID_NONE) return 0;
  uint32_t cnt = 0;
  if (p->fk_model) { /* SURVEY 8f row 3 */
    float sp[ORC_FK_MAX_SPHERES][3];
    orc_fk_model_points(p->fk_model, waypoint, p->d, sp);
    for (uint32_t q = 0; q < p->fk_model->num_spheres; ++q)
      cnt += (uint32_t)orc_grid_collides(g, sp[q][0], sp[q][1], sp[q][2],
                                         p->fk_model->spheres[q].radius + p->collision_distance);
    return cnt;
  }
  float pts[5][3];
  orc_fk_points(waypoint, p->d, pts);
  for (int l = 0; l < 5; ++l)
    cnt += (uint32_t)orc_grid_collides(g, pts[l][0], pts[l][1], pts[l][2], p->collision_distance);
  return cnt;
}

/* ------------------------------------------------------------------------
 * Stages 2-4 for one sample.
 * ---------------------------------------------------------------------- */
static float orc_tree32(const float *per_waypoint, uint32_t n) {
  float part[32];
  for (int l = 0; l < 32; ++l) part[l] = 0.0f;
  for (uint32_t w = 0; w < n; ++w) part[w & 31u] = part[w & 31u] + per_waypoint[w];
  for (int off = 16; off >= 1; off >>= 1) {
    float nxt[32];
    for (int l = 0; l < 32; ++l) nxt[l] = part[l] + part[l ^ off];
    memcpy(part, nxt, sizeof(part));
  }
  return part[0];
}

/* SEG4: per-waypoint sums are folded left to right inside segments of 4 consecutive
 * waypoints, then the segment partials are folded left to right.  This is the fixed order
 * of the CUDA kernels (one thread walks whole segments; any number of threads per sample
 * produces the same partials), replacing the reference's unordered atomicAdd (gt.cu:73-77). */
static float orc_seg4(const float *per_waypoint, uint32_t n) {
  float total = 0.0f;
  for (uint32_t w0 = 0; w0 < n; w0 += 4) {
    float part = 0.0f;
    for (uint32_t w = w0; w < n && w < w0 + 4; ++w) part = part + per_waypoint[w];
    total = total + part;
  }
  return total;
}

float orc_cost_sample(const orc_params *p, const orc_grid *g, const float *base,
                      const float *start, const float *noise, int32_t *collisions,
                      float *noisy_out, float *vel_out, float *acc_out, float *smooth_out,
                      int32_t *violated_out) {
  const uint32_t n = p->n, d = p->d;
  float xprev[8], vprev[8], x[8], v[8];
  float stack_sums[2 * 256]; /* per-waypoint sums: no heap traffic for the BASELINE shapes */
  float *acc2 = n <= 256 ? stack_sums : (float *)malloc(sizeof(float) * n * 2);
  float *lim2 = acc2 + n;
  int32_t coll = 0, violated = 0;
  float seq_sum = 0.0f, seq_lim = 0.0f; /* element-by-element order of gt.test.cu:187-195 */
  for (uint32_t j = 0; j < d; ++j) {
    xprev[j] = start[j]; /* gt.cu:285: initial_waypoint is the prior of waypoint 0 ... */
    vprev[j] = start[j]; /* ... for BOTH derivative passes (gt.test.cu:179-185) */
  }
  for (uint32_t w = 0; w < n; ++w) {
    const float *eps = noise ? noise + (size_t)(p->n_noise == 1 ? 0 : w) * d : NULL;
    float a2 = 0.0f, l2 = 0.0f;
    for (uint32_t j = 0; j < d; ++j) {
      x[j] = base[(size_t)w * d + j] + (eps ? eps[j] : 0.0f); /* gt.cu:53-57 */
      v[j] = (x[j] - xprev[j]) / p->delta_t;                  /* gt.cu:288 */
      float a = (v[j] - vprev[j]) / p->delta_t;               /* gt.cu:288 via :260-267 */
      float sq = a * a;                                       /* gt.cu:76 */
      a2 = a2 + sq;
      seq_sum = seq_sum + sq;
      if (noisy_out) noisy_out[(size_t)w * d + j] = x[j];
      if (vel_out) vel_out[(size_t)w * d + j] = v[j];
      if (acc_out) acc_out[(size_t)w * d + j] = a;
      if (p->limit_min && p->limit_max) {
        float over = x[j] - p->limit_max[j], under = p->limit_min[j] - x[j];
        float h = fmaxf(0.0f, fmaxf(over, under));
        if (h > 0.0f) violated = 1;
        float hs = h * h;
        l2 = l2 + hs;
        seq_lim = seq_lim + hs;
      }
    }
    acc2[w] = a2;
    lim2[w] = l2;
    coll += (int32_t)orc_num_collisions(p, g, x); /* gt.cu:86-103 */
    for (uint32_t j = 0; j < d; ++j) {
      xprev[j] = x[j];
      vprev[j] = v[j];
    }
  }
  float sum, lsum;
  if (p->sum_order == ORC_SUM_SEQUENTIAL) { /* gt.test.cu:187-195 */
    sum = seq_sum;
    lsum = seq_lim;
  } else if (p->sum_order == ORC_SUM_TREE32) {
    sum = orc_tree32(acc2, n);
    lsum = orc_tree32(lim2, n);
  } else if (p->sum_order == ORC_SUM_SEG4) {
    sum = orc_seg4(acc2, n);
    lsum = orc_seg4(lim2, n);
  } else { /* ORC_SUM_WAYPOINT: left fold of the per-waypoint sums, waypoint 0 first */
    sum = 0.0f;
    lsum = 0.0f;
    for (uint32_t w = 0; w < n; ++w) {
      sum = sum + acc2[w];
      lsum = lsum + lim2[w];
    }
  }
  float smooth = p->smooth_scale * sum; /* gt.cu:81,107 */
  float score = (float)coll + smooth;   /* gt.cu:100,108 */
  if (p->limit_weight != 0.0f && p->limit_min && p->limit_max) {
    float lc = p->limit_weight * lsum;
    score = score + lc;
  }
  if (acc2 != stack_sums) free(acc2);
  if (collisions) *collisions = coll;
  if (smooth_out) *smooth_out = smooth;
  if (violated_out) *violated_out = violated;
  return score;
}

static int32_t orc_round_score(float s) { /* gt.cu:109: int(ceilf(score)), saturating */
  return (s < 2147483648.0f) ? (int32_t)ceilf(s) : INT_MAX;
}

/* ---- threaded evaluation over samples (idiom of cdg.cu:88-110) ---------- */
typedef struct {
  const orc_params *p;
  const orc_grid *g;
  const float *base_traj, *start, *noise;
  uint32_t iteration;
  float *scores;
  int32_t *collisions;
  size_t s0, s1;
  size_t noise_m;      /* m of the noise buffer */
  uint64_t noise_t0;   /* global sample index of the buffer's first column */
} orc_job;

static void *orc_eval_range(void *arg) {
  orc_job *jb = (orc_job *)arg;
  const orc_params *p = jb->p;
  size_t nd = (size_t)p->n_noise * p->d;
  float tmp_small[8];
  float *tmp = nd <= 8 ? tmp_small : (float *)malloc(sizeof(float) * nd);
  for (size_t s = jb->s0; s < jb->s1; ++s) {
    uint32_t c = (uint32_t)(s / p->m);
    uint32_t t = (uint32_t)(s % p->m);
    uint64_t tg = p->sample_offset + t;
    const float *nz;
    if (jb->noise) {
      nz = jb->noise + ((size_t)c * jb->noise_m + (size_t)(tg - jb->noise_t0)) * nd;
    } else {
      for (uint32_t wn = 0; wn < p->n_noise; ++wn)
        for (uint32_t j = 0; j < p->d; ++j)
          tmp[(size_t)wn * p->d + j] = orc_noise_at(p, jb->iteration, c, tg, wn, j);
      nz = tmp;
    }
    int32_t coll;
    float sc = orc_cost_sample(p, jb->g, jb->base_traj + (size_t)c * p->n * p->d, jb->start, nz,
                               &coll, NULL, NULL, NULL, NULL, NULL);
    jb->scores[s] = sc;
    if (jb->collisions) jb->collisions[s] = coll;
  }
  if (tmp != tmp_small) free(tmp);
  return NULL;
}

/* Persistent worker pool: the threads are created once and reused by every evaluation (the
 * reference spawns its fill threads per call, cdg.cu:88-110, but it fills once per grid; the
 * planner loop evaluates once per ITERATION, so per-call pthread_create would dominate small
 * problems).  One evaluation at a time (mu_call); workers sleep on a condition variable. */
#define ORC_POOL_MAX 256
static struct {
  pthread_mutex_t mu, mu_call;
  pthread_cond_t cv_work, cv_done;
  pthread_t th[ORC_POOL_MAX];
  int nthreads;        /* workers alive */
  uint64_t generation; /* bumped per evaluation */
  orc_job *jobs;
  int njobs, next_job, pending;
} orc_pool = {PTHREAD_MUTEX_INITIALIZER, PTHREAD_MUTEX_INITIALIZER, PTHREAD_COND_INITIALIZER,
              PTHREAD_COND_INITIALIZER, {0}, 0, 0, NULL, 0, 0, 0};

static void *orc_pool_worker(void *arg) {
  (void)arg;
  pthread_mutex_lock(&orc_pool.mu);
  for (;;) {
    while (orc_pool.next_job >= orc_pool.njobs) pthread_cond_wait(&orc_pool.cv_work, &orc_pool.mu);
    orc_job *jb = &orc_pool.jobs[orc_pool.next_job++];
    pthread_mutex_unlock(&orc_pool.mu);
    orc_eval_range(jb);
    pthread_mutex_lock(&orc_pool.mu);
    if (--orc_pool.pending == 0) pthread_cond_signal(&orc_pool.cv_done);
  }
  return NULL;
}

static void orc_eval_all(const orc_params *p, const orc_grid *g, const float *base_traj,
                         const float *start, const float *noise, size_t noise_m,
                         uint64_t noise_t0, uint32_t iteration, float *scores,
                         int32_t *collisions, int num_threads) {
  size_t S = (size_t)p->k * p->m;
  if (num_threads < 1) num_threads = 1;
  if (num_threads > ORC_POOL_MAX) num_threads = ORC_POOL_MAX;
  if ((size_t)num_threads > S) num_threads = (int)(S ? S : 1);
  orc_job jobs[ORC_POOL_MAX];
  size_t per = (S + (size_t)num_threads - 1) / (size_t)num_threads;
  for (int i = 0; i < num_threads; ++i) {
    orc_job jb = {p, g, base_traj, start, noise, iteration, scores, collisions,
                  0, 0, noise_m, noise_t0};
    jb.s0 = per * (size_t)i < S ? per * (size_t)i : S;
    jb.s1 = per * (size_t)(i + 1) < S ? per * (size_t)(i + 1) : S;
    jobs[i] = jb;
  }
  if (num_threads == 1) {
    orc_eval_range(&jobs[0]);
    return;
  }
  /* the calling thread takes job 0, the pool the rest */
  pthread_mutex_lock(&orc_pool.mu_call);
  pthread_mutex_lock(&orc_pool.mu);
  while (orc_pool.nthreads < num_threads - 1) {
    if (pthread_create(&orc_pool.th[orc_pool.nthreads], NULL, orc_pool_worker, NULL) != 0) break;
    pthread_detach(orc_pool.th[orc_pool.nthreads]);
    orc_pool.nthreads++;
  }
  const int have_pool = orc_pool.nthreads > 0;
  if (have_pool) {
    orc_pool.jobs = jobs;
    orc_pool.next_job = 1;
    orc_pool.pending = num_threads - 1;
    orc_pool.njobs = num_threads;
    orc_pool.generation++;
    pthread_cond_broadcast(&orc_pool.cv_work);
  }
  pthread_mutex_unlock(&orc_pool.mu);
  orc_eval_range(&jobs[0]);
  if (have_pool) {
    pthread_mutex_lock(&orc_pool.mu);
    while (orc_pool.pending > 0) pthread_cond_wait(&orc_pool.cv_done, &orc_pool.mu);
    orc_pool.njobs = 0;
    orc_pool.next_job = 0;
    orc_pool.jobs = NULL;
    pthread_mutex_unlock(&orc_pool.mu);
  } else {
    for (int i = 1; i < num_threads; ++i) orc_eval_range(&jobs[i]);
  }
  pthread_mutex_unlock(&orc_pool.mu_call);
}

int64_t orc_cost_eval(const orc_params *p, const orc_grid *g, const float *base_traj,
                      const float *start, const float *noise, uint32_t iteration,
                      float *scores, int32_t *collisions, float *min_score, int num_threads) {
  size_t S = (size_t)p->k * p->m;
  float *sc = scores ? scores : (float *)malloc(sizeof(float) * (S ? S : 1));
  orc_eval_all(p, g, base_traj, start, noise, p->m, p->sample_offset, iteration, sc, collisions,
               num_threads);
  /* ps.cu:113-122: strict <, first lowest wins, starts at +inf */
  uint64_t mg = p->m_global ? p->m_global : p->m;
  float lowest = INFINITY;
  int64_t best = -1;
  for (size_t s = 0; s < S; ++s) {
    if (sc[s] < lowest) {
      lowest = sc[s];
      best = (int64_t)((s / p->m) * mg + p->sample_offset + (s % p->m));
    }
  }
  if (min_score) *min_score = lowest;
  if (!scores) free(sc);
  return best;
}

void orc_plan_lerp(const float *start, const float *goal, uint32_t n, uint32_t d, float *out) {
  /* ps.cu:50-52: step_sizes[dim] = (goal - current) / num_waypoints; the write
   * of `result` is missing in the reference. */
  for (uint32_t j = 0; j < d; ++j) {
    float step = (goal[j] - start[j]) / (float)n;
    for (uint32_t w = 0; w < n; ++w) out[(size_t)w * d + j] = start[j] + (float)(w + 1) * step;
  }
}

static uint32_t orc_key_hi(const orc_params *p, float score) {
  if (p->select_mode == ORC_SELECT_CEIL_INT) return (uint32_t)orc_round_score(score);
  if (score != score) return 0xFFFFFFFFu;
  uint32_t b;
  memcpy(&b, &score, 4);
  return b;
}

void orc_plan(const orc_params *p, const orc_grid *g, const float *start, const float *init_traj,
              uint32_t iters, const float *noise, float *best_traj, float *best_vel,
              float *best_score, int64_t *best_climber, float *climber_traj,
              float *climber_scores, int32_t *climber_rounded, uint8_t *found_better,
              int num_threads) {
  const uint32_t k = p->k, m = p->m, n = p->n, d = p->d;
  const size_t nd = (size_t)n * d, nnd = (size_t)p->n_noise * d;
  float *traj = (float *)malloc(sizeof(float) * k * nd);
  uint32_t *cur_hi = (uint32_t *)malloc(sizeof(uint32_t) * k);
  uint8_t *fb = (uint8_t *)calloc(k, 1);
  float *scores = (float *)malloc(sizeof(float) * (size_t)k * m);
  float *tmpn = (float *)malloc(sizeof(float) * nnd);
  for (uint32_t c = 0; c < k; ++c) memcpy(traj + c * nd, init_traj, sizeof(float) * nd);
  /* gt.cu:180-191: score the unperturbed trajectory first (noise == 0) */
  for (uint32_t c = 0; c < k; ++c) {
    float s0 = orc_cost_sample(p, g, traj + c * nd, start, NULL, NULL, NULL, NULL, NULL, NULL, NULL);
    cur_hi[c] = orc_key_hi(p, s0);
  }
  uint64_t mg = p->m_global ? p->m_global : m;
  for (uint32_t it = 0; it < iters; ++it) {
    const float *nz_it = noise ? noise + (size_t)it * k * mg * nnd : NULL;
    orc_eval_all(p, g, traj, start, nz_it, (size_t)mg, 0, it, scores, NULL, num_threads);
    if (p->update_mode == ORC_UPDATE_WEIGHTED) {
      /* gt.cu:242-251 (commented out in the reference): u = sum_t (best/ceil(score_t)) noise_t,
       * traj[w][j] -= gain * u[j]; n_noise == 1 */
      float *nzc = nz_it ? NULL : (float *)malloc(sizeof(float) * (size_t)m * d);
      for (uint32_t c = 0; c < k; ++c) {
        int32_t best = INT_MAX;
        for (uint32_t t = 0; t < m; ++t) {
          int32_t r = orc_round_score(scores[(size_t)c * m + t]);
          if (r < best) best = r;
        }
        const float *nz = nz_it ? nz_it + ((size_t)c * mg + p->sample_offset) * d : nzc;
        if (!nz_it)
          for (uint32_t t = 0; t < m; ++t)
            for (uint32_t j = 0; j < d; ++j)
              nzc[(size_t)t * d + j] = orc_noise_at(p, it, c, p->sample_offset + t, 0, j);
        int64_t sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        orc_update_sums(p, best, scores + (size_t)c * m, nz, m, sums);
        for (uint32_t j = 0; j < d; ++j) {
          float delta = p->update_gain * orc_update_from_sum(sums[j], sums[7]);
          for (uint32_t w = 0; w < n; ++w)
            traj[c * nd + (size_t)w * d + j] = traj[c * nd + (size_t)w * d + j] - delta;
        }
        fb[c] = 1;
      }
      free(nzc);
      continue;
    }
    for (uint32_t c = 0; c < k; ++c) {
      /* gt.cu:109-111 atomicMin over the block, then :225-241 keep-best with
       * the lowest sample index winning ties */
      uint64_t best_key = UINT64_MAX;
      for (uint32_t t = 0; t < m; ++t) {
        uint64_t key = ((uint64_t)orc_key_hi(p, scores[(size_t)c * m + t]) << 32) |
                       (uint32_t)(p->sample_offset + t);
        if (key < best_key) best_key = key;
      }
      uint32_t hi = (uint32_t)(best_key >> 32);
      uint64_t tg = (uint32_t)best_key;
      if (hi < cur_hi[c]) {
        const float *nz;
        if (nz_it) {
          nz = nz_it + ((size_t)c * mg + tg) * nnd;
        } else {
          for (uint32_t wn = 0; wn < p->n_noise; ++wn)
            for (uint32_t j = 0; j < d; ++j) tmpn[wn * d + j] = orc_noise_at(p, it, c, tg, wn, j);
          nz = tmpn;
        }
        for (uint32_t w = 0; w < n; ++w)
          for (uint32_t j = 0; j < d; ++j)
            traj[c * nd + (size_t)w * d + j] =
                traj[c * nd + (size_t)w * d + j] + nz[(size_t)(p->n_noise == 1 ? 0 : w) * d + j];
        cur_hi[c] = hi;
        fb[c] = 1;
      }
    }
  }
  /* final: float score of every climber's kept trajectory, then ps.cu:113-122 */
  float lowest = INFINITY;
  int64_t bc = -1;
  for (uint32_t c = 0; c < k; ++c) {
    float s = orc_cost_sample(p, g, traj + c * nd, start, NULL, NULL, NULL, NULL, NULL, NULL, NULL);
    if (climber_scores) climber_scores[c] = s;
    if (climber_rounded) climber_rounded[c] = orc_round_score(s);
    if (s < lowest) {
      lowest = s;
      bc = c;
    }
  }
  uint32_t sel = bc < 0 ? 0 : (uint32_t)bc;
  if (best_traj) memcpy(best_traj, traj + sel * nd, sizeof(float) * nd);
  if (best_vel)
    for (uint32_t w = 0; w < n; ++w)
      for (uint32_t j = 0; j < d; ++j) {
        float prior = w ? traj[sel * nd + (size_t)(w - 1) * d + j] : start[j];
        best_vel[(size_t)w * d + j] = (traj[sel * nd + (size_t)w * d + j] - prior) / p->delta_t;
      }
  if (best_score) *best_score = bc < 0 ? INFINITY : lowest;
  if (best_climber) *best_climber = (bc >= 0 && fb[bc]) ? bc : -1;
  if (climber_traj) memcpy(climber_traj, traj, sizeof(float) * k * nd);
  if (found_better) memcpy(found_better, fb, k);
  free(traj);
  free(cur_hi);
  free(fb);
  free(scores);
  free(tmpn);
}

void orc_update_vector(const orc_params *p, const float *scores, const float *noise,
                       float *update) {
  /* gt.test.cu:200-208 (host statement of gt.cu:117-133); n_noise == 1 */
  size_t S = (size_t)p->k * p->m;
  float best = scores[0];
  for (size_t s = 1; s < S; ++s)
    if (ceilf(scores[s]) < ceilf(best)) best = scores[s]; /* gt.test.cu:201 */
  for (uint32_t c = 0; c < p->k; ++c)
    for (uint32_t j = 0; j < p->d; ++j) update[(size_t)c * p->d + j] = 0.0f;
  for (uint32_t c = 0; c < p->k; ++c)
    for (uint32_t t = 0; t < p->m; ++t)
      for (uint32_t j = 0; j < p->d; ++j) {
        float wgt = best / ceilf(scores[(size_t)c * p->m + t]);
        float term = wgt * noise[((size_t)c * p->m + t) * p->d + j];
        update[(size_t)c * p->d + j] = update[(size_t)c * p->d + j] + term;
      }
}

/* ---- exact Euclidean distance transform ------------------------------------------------ */
#define ORC_EDT_INF (1 << 29)
static int64_t orc_floordiv(int64_t a, int64_t b) { /* b > 0 */
  int64_t q = a / b;
  if ((a % b != 0) && (a < 0)) --q;
  return q;
}
/* one line of the lower-envelope scan: g[] squared distances so far, out[] their 1-D transform */
static void orc_edt_line(const int32_t *g, int32_t *out, int32_t n, int32_t *s, int32_t *t) {
  int32_t q = 0;
  s[0] = 0;
  t[0] = 0;
#define ORC_F(x, i) ((int64_t)((x) - (i)) * ((x) - (i)) + g[i])
  for (int32_t u = 1; u < n; ++u) {
    while (q >= 0 && ORC_F(t[q], s[q]) > ORC_F(t[q], u)) --q;
    if (q < 0) {
      q = 0;
      s[0] = u;
    } else {
      int64_t i = s[q];
      int64_t w = 1 + orc_floordiv((int64_t)u * u - i * i + g[u] - g[i], 2 * (u - i));
      if (w < n) {
        ++q;
        s[q] = u;
        t[q] = (int32_t)w;
      }
    }
  }
  for (int32_t u = n - 1; u >= 0; --u) {
    int64_t f = ORC_F(u, s[q]);
    out[u] = f >= ORC_EDT_INF ? ORC_EDT_INF : (int32_t)f;
    if (u == t[q]) --q;
  }
#undef ORC_F
}

void orc_distance_field(const orc_grid *occ, float *out) {
  const int32_t nx = (int32_t)occ->nx, ny = (int32_t)occ->ny, nz = (int32_t)occ->nz;
  const size_t cells = (size_t)nx * ny * nz;
  const uint8_t *c = (const uint8_t *)occ->cells;
  int32_t *a = (int32_t *)malloc(sizeof(int32_t) * cells), *b = (int32_t *)malloc(sizeof(int32_t) * cells);
  int32_t nmax = nx > ny ? nx : ny;
  if (nz > nmax) nmax = nz;
  int32_t *g = (int32_t *)malloc(sizeof(int32_t) * nmax * 4), *o = g + nmax, *s = o + nmax, *t = s + nmax;
  /* along x: distance to the nearest occupied voxel of the row, squared */
  for (size_t row = 0; row < (size_t)ny * nz; ++row) {
    const uint8_t *r = c + row * nx;
    int32_t *d = a + row * nx;
    int32_t last = -ORC_EDT_INF;
    for (int32_t x = 0; x < nx; ++x) {
      if (r[x]) last = x;
      d[x] = last < 0 ? ORC_EDT_INF : x - last;
    }
    last = ORC_EDT_INF;
    for (int32_t x = nx - 1; x >= 0; --x) {
      if (r[x]) last = x;
      int32_t dr = last >= ORC_EDT_INF ? ORC_EDT_INF : last - x;
      int32_t m = d[x] < dr ? d[x] : dr;
      d[x] = m >= ORC_EDT_INF ? ORC_EDT_INF : m * m;
    }
  }
  /* along y, then along z */
  for (int32_t z = 0; z < nz; ++z)
    for (int32_t x = 0; x < nx; ++x) {
      for (int32_t y = 0; y < ny; ++y) g[y] = a[((size_t)z * ny + y) * nx + x];
      orc_edt_line(g, o, ny, s, t);
      for (int32_t y = 0; y < ny; ++y) b[((size_t)z * ny + y) * nx + x] = o[y];
    }
  const float res = (float)occ->resolution;
  for (int32_t y = 0; y < ny; ++y)
    for (int32_t x = 0; x < nx; ++x) {
      for (int32_t z = 0; z < nz; ++z) g[z] = b[((size_t)z * ny + y) * nx + x];
      orc_edt_line(g, o, nz, s, t);
      for (int32_t z = 0; z < nz; ++z)
        out[((size_t)z * ny + y) * nx + x] = o[z] >= ORC_EDT_INF ? INFINITY : sqrtf((float)o[z]) * res;
    }
  free(a);
  free(b);
  free(g);
}

void orc_distance_field_brute(const orc_grid *occ, float *out) {
  const int32_t nx = (int32_t)occ->nx, ny = (int32_t)occ->ny, nz = (int32_t)occ->nz;
  const uint8_t *c = (const uint8_t *)occ->cells;
  const float res = (float)occ->resolution;
  for (int32_t z = 0; z < nz; ++z)
    for (int32_t y = 0; y < ny; ++y)
      for (int32_t x = 0; x < nx; ++x) {
        int64_t best = -1;
        for (int32_t k = 0; k < nz; ++k)
          for (int32_t j = 0; j < ny; ++j)
            for (int32_t i = 0; i < nx; ++i)
              if (c[((size_t)k * ny + j) * nx + i]) {
                int64_t d = (int64_t)(x - i) * (x - i) + (int64_t)(y - j) * (y - j) + (int64_t)(z - k) * (z - k);
                if (best < 0 || d < best) best = d;
              }
        out[((size_t)z * ny + y) * nx + x] = best < 0 ? INFINITY : sqrtf((float)best) * res;
      }
}

uint64_t orc_depth_scatter(orc_grid *g, const uint16_t *depth, uint32_t cams, uint32_t w, uint32_t h,
                           const double *intrinsics, const float *extrinsics, double crop_size,
                           int8_t val, int clear_first) {
  size_t cells = (size_t)g->nx * g->ny * g->nz;
  if (clear_first) memset(g->cells, 0, cells);
  const float half = (float)(crop_size / 2.0), full = (float)crop_size; /* map.cpp:262-263 */
  uint64_t kept = 0;
  for (uint32_t c = 0; c < cams; ++c) {
    const double unit_scaling = (double)0.001f; /* DepthTraits<uint16_t>::toMeters(1) */
    const float center_x = (float)intrinsics[4 * c + 2], center_y = (float)intrinsics[4 * c + 3];
    const float constant_x = (float)(unit_scaling / intrinsics[4 * c]);     /* map.cpp:101-103 */
    const float constant_y = (float)(unit_scaling / intrinsics[4 * c + 1]);
    const float *r = extrinsics + 12 * c, *t = r + 9;
    for (uint32_t v = 0; v < h; ++v)
      for (uint32_t u = 0; u < w; ++u) {
        uint16_t d = depth[((size_t)c * h + v) * w + u];
        if (d == 0) continue; /* map.cpp:117-126, range_max == 0 */
        float df = (float)d;
        float x = ((float)u - center_x) * df; x = x * constant_x; /* map.cpp:130 */
        float y = ((float)v - center_y) * df; y = y * constant_y; /* :131 */
        float z = df * 0.001f;                                    /* :132 */
        float q[3];
        for (int a = 0; a < 3; ++a) {
          float s = r[3 * a] * x;
          float s1 = r[3 * a + 1] * y;
          float s2 = r[3 * a + 2] * z;
          s = s + s1;
          s = s + s2;
          q[a] = s + t[a];
        }
        if (!(q[0] >= -half && q[0] <= half && q[1] >= -half && q[1] <= half && q[2] >= 0.0f &&
              q[2] <= full))
          continue;
        kept++;
        int64_t idx = orc_grid_index(g, q[0], q[1], q[2]);
        if (idx >= 0) ((uint8_t *)g->cells)[idx] = (uint8_t)val;
      }
  }
  return kept;
}

void orc_update_sums(const orc_params *p, int32_t best_rounded, const float *scores,
                     const float *noise, uint32_t count, int64_t sums[8]) {
  /* gt.cu:128-132: weight = best_score / ceilf(our_score); term = weight * noise */
  const float best = (float)best_rounded;
  for (uint32_t t = 0; t < count; ++t) {
    float wgt = best / ceilf(scores[t]);
    for (uint32_t j = 0; j < p->d; ++j) {
      float term = wgt * noise[(size_t)t * p->d + j];
      if (!(fabsf(term) < 4096.0f)) {
        sums[7] += 1;
        continue;
      }
      sums[j] += (int64_t)llrintf(term * 268435456.0f); /* exact scaling by 2^28, RN-even */
    }
  }
}

float orc_update_from_sum(int64_t sum, int64_t poisoned) {
  if (poisoned) return NAN;
  return (float)((double)sum * 0x1p-28);
}

void orc_build_message(const float *cur, const float *goal, const float *traj, const float *vel,
                       uint32_t n, uint32_t d, float delta_t, float *positions,
                       float *velocities, float *time_from_start) {
  /* ps.cu:179-191: point 0 = current state, velocity towards waypoint 0 */
  for (uint32_t j = 0; j < d; ++j) {
    positions[j] = cur[j];
    velocities[j] = (traj[j] - cur[j]) / delta_t;
  }
  time_from_start[0] = 0.0f;
  /* ps.cu:192-202: look-forward velocities; the read of waypoint n runs past
   * the end in the reference -> 0 for the last waypoint */
  for (uint32_t i = 0; i < n; ++i) {
    for (uint32_t j = 0; j < d; ++j) {
      positions[(size_t)(i + 1) * d + j] = traj[(size_t)i * d + j];
      velocities[(size_t)(i + 1) * d + j] = (i + 1 < n) ? vel[(size_t)(i + 1) * d + j] : 0.0f;
    }
    time_from_start[i + 1] = delta_t * (float)i;
  }
  /* ps.cu:204-208: goal with zero velocity */
  for (uint32_t j = 0; j < d; ++j) {
    positions[(size_t)(n + 1) * d + j] = goal[j];
    velocities[(size_t)(n + 1) * d + j] = 0.0f;
  }
  /* the reference reuses the moved-from `point`, so the goal keeps the stamp of
   * the last waypoint (ps.cu:200,204-208) */
  time_from_start[n + 1] = n ? delta_t * (float)(n - 1) : 0.0f;
}
