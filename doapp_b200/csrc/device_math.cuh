// device_math.cuh — deterministic elementary functions and Philox for the
// sm_100a kernels.  Written from the specification in DESIGN.md ("deterministic
// math spec"); the oracle holds an independent C restatement of the same spec
// and tests/test_parity_gpu.py compares the two bit for bit.
//
// The reference calls std::sin / std::cos on the device
// (perturbator/src/generate_trajectories.cu:304-305) and curand_normal
// (:297-299); both are replaced here: sin/cos because CUDA's and glibc's
// results differ in the last ulp (which flips voxel indices), cuRAND by the
// counter-based Philox4x32-10 the north star asks for.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define DM_HD __host__ __device__ __forceinline__
#else
#define DM_HD inline
#endif

namespace dm {

// single-rounding helpers: immune to -fmad / host FP contraction settings
DM_HD float fma_(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return fmaf(a, b, c);
#endif
}
DM_HD float mul_(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fmul_rn(a, b);
#else
  volatile float r = a * b;
  return r;
#endif
}
DM_HD float add_(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fadd_rn(a, b);
#else
  volatile float r = a + b;
  return r;
#endif
}
DM_HD float sub_(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fsub_rn(a, b);
#else
  volatile float r = a - b;
  return r;
#endif
}
DM_HD float div_(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdiv_rn(a, b);
#else
  volatile float r = a / b;
  return r;
#endif
}

// ---------------------------------------------------------------------------
// Division by a loop-invariant divisor b with a precomputed r = RN(1/b):
//   q0 = RN(a*r); e = fma(-b, q0, a) (exact); q = fma(e, r, q0)
// (Markstein).  For b, r normal and |a| in [2^-60, 2^60] (or a == 0) no
// intermediate is subnormal or overflows and q == RN(a/b); the C ABI verifies
// this EXHAUSTIVELY on the device for each divisor it is given
// (doapp_capi.cu: verify_fast_div) and the kernels fall back to the IEEE
// divide outside that interval, so every quotient is bit-identical to a / b.
struct FastDiv {
  float b, r;
  float lo, hi; // |a| in [lo, hi) (or a == 0) takes the fast path; lo = +inf disables it
};
constexpr float kFastDivLo = 0x1p-98f, kFastDivHi = 0x1p+60f;
// The kernels do not test the 10 dividends of a waypoint but its 5 joint angles: if every
// angle x of three consecutive waypoints is 0 or kFineLo <= |x| < kFineHi, and the divisor dt
// lies in [kFastDtLo, kFastDtHi], then x - xprev is 0 or a multiple of 2^-59 below 2^18, the
// velocity v is 0 or >= 2^-59/dt, and v - vprev is 0 or a multiple of ulp(2^-59/dt) >=
// 2^-84/dt >= 2^-94: all inside [kFastDivLo, kFastDivHi).  kFineHi is also the range of
// sincos_small.
constexpr float kFineLo = 0x1p-36f, kFineHi = 0x1p+17f;
constexpr float kFastDtLo = 0x1p-20f, kFastDtHi = 0x1p+10f;
DM_HD bool fine_angle(float x) {
  const float a = fabsf(x);
  return (a == 0.0f) | ((a >= kFineLo) & (a < kFineHi));
}

DM_HD float div_fast_unguarded(float a, FastDiv d) {
  float q0 = mul_(a, d.r);
  float e = fma_(-d.b, q0, a);
  return fma_(e, d.r, q0);
}
DM_HD float div_fast(float a, FastDiv d) {
  float q = div_fast_unguarded(a, d);
  float aa = fabsf(a);
  if (!(aa == 0.0f || (aa >= d.lo && aa < d.hi))) q = div_(a, d.b);
  return q;
}

constexpr float kTwoOverPi = 0x1.45f306p-1f;
constexpr float kP1 = 0x1.921fb6p+0f;   // RN(pi/2)
constexpr float kP2 = -0x1.777a5cp-25f; // RN(pi/2 - P1)
constexpr float kP3 = -0x1.ee59dap-50f; // RN(pi/2 - P1 - P2)
constexpr double kTwoOverPiD = 0x1.45f306dc9c883p-1;
constexpr double kTwoPiD = 0x1.921fb54442d18p+2;
constexpr double kP1D = 0x1.921fb54442d18p+0;
constexpr double kP2D = 0x1.1a62633145c07p-54;
constexpr float kS1 = -1.6666654611e-1f, kS2 = 8.3321608736e-3f, kS3 = -1.9515295891e-4f;
constexpr float kC1 = 4.166664568298827e-2f, kC2 = -1.388731625493765e-3f,
                kC3 = 2.443315711809948e-5f;

// Cold path: |x| >= 2^17, inf or nan.
#if defined(__CUDACC__)
inline __host__ __device__ __noinline__
#else
inline
#endif
    void
    sincos_slow(float x, float *r_out, int *q_out) {
  double xd = fmod((double)x, kTwoPiD);
  double kd = rint(xd * kTwoOverPiD);
  double rd = fma(-kd, kP1D, xd);
  rd = fma(-kd, kP2D, rd);
  *r_out = (float)rd;
  *q_out = (int)kd & 3;
}

DM_HD uint32_t f2u_(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
#endif
}
DM_HD float u2f_(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// polynomial core + quadrant fix-up shared by both argument ranges; q = k mod 4
DM_HD void sincos_core(float r, uint32_t q, float *s, float *c) {
  float z = mul_(r, r);
  float ps = fma_(kS3, z, kS2);
  ps = fma_(ps, z, kS1);
  ps = mul_(ps, z);
  float sin_r = fma_(ps, r, r);
  float pc = fma_(kC3, z, kC2);
  pc = fma_(pc, z, kC1);
  pc = fma_(pc, z, -0.5f);
  float cos_r = fma_(pc, z, 1.0f);
  // (s,c) = (sin,cos), (cos,-sin), (-sin,-cos), (-cos,sin) for q = 0..3, as sign-bit flips
  const bool swap = (q & 1u) != 0;
  const float sa = swap ? cos_r : sin_r, ca = swap ? sin_r : cos_r;
  const uint32_t qs = q << 30; // bit 31 = bit 1 of q
  *s = u2f_(f2u_(sa) ^ (qs & 0x80000000u));
  *c = u2f_(f2u_(ca) ^ ((qs + 0x40000000u) & 0x80000000u)); // bit 1 of q + 1
}

// |x| < 2^17 only (the caller checks): straight-line code, no conversions.  kf = rintf(t) is
// taken with the 1.5*2^23 trick (t + M rounds to an integer, ties to even, exactly as rintf
// does for |t| < 2^22) and k mod 4 is read from the low mantissa bits of t + M.
DM_HD void sincos_small(float x, float *s, float *c) {
  const float kMagic = 12582912.0f; // 1.5 * 2^23
  float t = mul_(x, kTwoOverPi);
  float tm = add_(t, kMagic);
  float kf = sub_(tm, kMagic);
  float r = fma_(-kf, kP1, x);
  r = fma_(-kf, kP2, r);
  r = fma_(-kf, kP3, r);
  sincos_core(r, f2u_(tm) & 3u, s, c);
}

DM_HD void sincos(float x, float *s, float *c) {
  if (fabsf(x) < 131072.0f) {
    sincos_small(x, s, c);
  } else if (x != x || fabsf(x) == INFINITY) {
    *s = NAN;
    *c = NAN;
  } else {
    float r;
    int q;
    sincos_slow(x, &r, &q);
    sincos_core(r, (uint32_t)q & 3u, s, c);
  }
}

#if defined(__CUDACC__)
// ---------------------------------------------------------------------------
// Packed fp32x2 arithmetic (sm_100a FFMA2 / FADD2 / FMUL2): one issue slot carries the same
// operation for two samples.  Every operation is an IEEE round-to-nearest fma per half, so
// the results equal the scalar helpers above bit for bit.  add / mul are written as
// fma(a, 1, b) / fma(a, b, -0) (exact identities, ptxas emits FADD2 / FMUL2).
//
// HAZARD: ptxas contracts a packed multiply followed by a packed add of its result into one
// FFMA2 even under --fmad=false (observed with CUDA 12.9).  Wherever the specification
// rounds the product before the addition (acc*acc summed, t = x*(2/pi) then t + magic), the
// product must therefore be formed with the scalar __fmul_rn per half (f2_mul_rounded), never
// with f2_mul.
struct f2 {
  unsigned long long v;
};
__device__ __forceinline__ f2 f2_pack(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(f2 a, float *lo, float *hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(*lo), "=f"(*hi) : "l"(a.v));
}
__device__ __forceinline__ float f2_lo(f2 a) {
  float lo, hi;
  f2_unpack(a, &lo, &hi);
  return lo;
}
__device__ __forceinline__ float f2_hi(f2 a) {
  float lo, hi;
  f2_unpack(a, &lo, &hi);
  return hi;
}
__device__ __forceinline__ f2 f2_splat(float c) { return f2_pack(c, c); }
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
__device__ __forceinline__ f2 f2_neg(f2 a) { // folds into an operand modifier
  float lo, hi;
  f2_unpack(a, &lo, &hi);
  return f2_pack(-lo, -hi);
}
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) { return f2_fma(a, f2_splat(1.0f), b); }
__device__ __forceinline__ f2 f2_sub(f2 a, f2 b) { return f2_fma(b, f2_splat(-1.0f), a); }
// product consumed as an fma operand or converted, never by an add/sub (see HAZARD)
__device__ __forceinline__ f2 f2_mul(f2 a, f2 b) { return f2_fma(a, b, f2_splat(-0.0f)); }
// product that the specification rounds before adding it to something
__device__ __forceinline__ f2 f2_mul_rounded(f2 a, f2 b) {
  float al, ah, bl, bh;
  f2_unpack(a, &al, &ah);
  f2_unpack(b, &bl, &bh);
  return f2_pack(__fmul_rn(al, bl), __fmul_rn(ah, bh));
}
__device__ __forceinline__ f2 f2_div_fast_unguarded(f2 a, FastDiv d) {
  const f2 r = f2_splat(d.r);
  f2 q0 = f2_mul(a, r);
  f2 e = f2_fma(f2_splat(-d.b), q0, a);
  return f2_fma(e, r, q0);
}
// sincos_small on two angles; quadrant fix-up per half as in sincos_core
__device__ __forceinline__ void f2_sincos_small(f2 x, f2 *s, f2 *c) {
  const float kMagic = 12582912.0f;
  f2 t = f2_mul_rounded(x, f2_splat(kTwoOverPi));
  f2 tm = f2_add(t, f2_splat(kMagic));
  f2 kf = f2_add(tm, f2_splat(-kMagic));
  f2 r = f2_fma(kf, f2_splat(-kP1), x); // fma(-kf, P, x) == fma(kf, -P, x)
  r = f2_fma(kf, f2_splat(-kP2), r);
  r = f2_fma(kf, f2_splat(-kP3), r);
  f2 z = f2_mul(r, r);
  f2 ps = f2_fma(f2_splat(kS3), z, f2_splat(kS2));
  ps = f2_fma(ps, z, f2_splat(kS1));
  ps = f2_mul(ps, z);
  f2 sin_r = f2_fma(ps, r, r);
  f2 pc = f2_fma(f2_splat(kC3), z, f2_splat(kC2));
  pc = f2_fma(pc, z, f2_splat(kC1));
  pc = f2_fma(pc, z, f2_splat(-0.5f));
  f2 cos_r = f2_fma(pc, z, f2_splat(1.0f));
  float sl, sh, cl, ch, tl, th;
  f2_unpack(sin_r, &sl, &sh);
  f2_unpack(cos_r, &cl, &ch);
  f2_unpack(tm, &tl, &th);
  auto fix = [](float sin_v, float cos_v, uint32_t q, float *so, float *co) {
    const bool swap = (q & 1u) != 0;
    const float sa = swap ? cos_v : sin_v, ca = swap ? sin_v : cos_v;
    const uint32_t qs = q << 30;
    *so = __uint_as_float(__float_as_uint(sa) ^ (qs & 0x80000000u));
    *co = __uint_as_float(__float_as_uint(ca) ^ ((qs + 0x40000000u) & 0x80000000u));
  };
  float s0, s1, c0, c1;
  fix(sl, cl, __float_as_uint(tl) & 3u, &s0, &c0);
  fix(sh, ch, __float_as_uint(th) & 3u, &s1, &c1);
  *s = f2_pack(s0, s1);
  *c = f2_pack(c0, c1);
}
#endif // __CUDACC__

// natural log on (0, 1]
DM_HD float log_(float x) {
#if defined(__CUDA_ARCH__)
  uint32_t ix = __float_as_uint(x);
#else
  uint32_t ix;
  memcpy(&ix, &x, 4);
#endif
  ix += 0x3f800000u - 0x3f3504f3u;
  int e = (int)(ix >> 23) - 127;
  ix = (ix & 0x007fffffu) + 0x3f3504f3u;
#if defined(__CUDA_ARCH__)
  float mant = __uint_as_float(ix);
#else
  float mant;
  memcpy(&mant, &ix, 4);
#endif
  float f = sub_(mant, 1.0f);
  float z = mul_(f, f);
  float p = 7.0376836292e-2f;
  p = fma_(p, f, -1.1514610310e-1f);
  p = fma_(p, f, 1.1676998740e-1f);
  p = fma_(p, f, -1.2420140846e-1f);
  p = fma_(p, f, 1.4249322787e-1f);
  p = fma_(p, f, -1.6668057665e-1f);
  p = fma_(p, f, 2.0000714765e-1f);
  p = fma_(p, f, -2.4999993993e-1f);
  p = fma_(p, f, 3.3333331174e-1f);
  float y = mul_(mul_(f, z), p);
  float fe = (float)e;
  y = fma_(fe, -2.12194440e-4f, y);
  y = fma_(-0.5f, z, y);
  float r = add_(f, y);
  r = fma_(fe, 0.693359375f, r);
  return r;
}

struct u4 {
  uint32_t x, y, z, w;
};

DM_HD u4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                       uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
#if defined(__CUDA_ARCH__)
    uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
#else
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t h0 = (uint32_t)(p0 >> 32), l0 = (uint32_t)p0;
    uint32_t h1 = (uint32_t)(p1 >> 32), l1 = (uint32_t)p1;
#endif
    uint32_t n0 = h1 ^ c1 ^ k0;
    uint32_t n2 = h0 ^ c3 ^ k1;
    c0 = n0;
    c1 = l1;
    c2 = n2;
    c3 = l0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return u4{c0, c1, c2, c3};
}

// one N(0, sigma^2) draw from two Philox words: cos branch for even joints
DM_HD float box_muller(uint32_t a, uint32_t b, bool want_sin, float sigma) {
  float u1 = mul_((float)((a >> 8) + 1u), 0x1p-24f);
  float u2 = mul_((float)(b >> 8), 0x1p-24f);
  float rad = sqrtf(mul_(-2.0f, log_(u1)));
  float s, c;
  sincos(mul_(u2, 0x1.921fb6p+2f), &s, &c);
  return mul_(sigma, mul_(rad, want_sin ? s : c));
}

// noise element (climber c, global sample t, noise waypoint wn, joint j)
DM_HD float noise_at(uint32_t seed_lo, uint32_t seed_hi, float sigma, uint32_t iteration,
                     uint32_t c, uint32_t t_global, uint32_t wn, uint32_t j) {
  u4 r = philox4x32_10(t_global, c, wn, 2u * iteration + (j >> 2), seed_lo, seed_hi);
  uint32_t a = (j & 2u) ? r.z : r.x;
  uint32_t b = (j & 2u) ? r.w : r.y;
  return box_muller(a, b, (j & 1u) != 0, sigma);
}

// all D noise elements of one (sample, noise waypoint): joints 0-3 share one Philox
// call, joint 4 takes the second (same values as noise_at element by element)
template <int D>
DM_HD void noise_vec(uint32_t seed_lo, uint32_t seed_hi, float sigma, uint32_t iteration,
                     uint32_t c, uint32_t t_global, uint32_t wn, float (&out)[D]) {
  u4 r = philox4x32_10(t_global, c, wn, 2u * iteration, seed_lo, seed_hi);
  out[0] = box_muller(r.x, r.y, false, sigma);
  out[1] = box_muller(r.x, r.y, true, sigma);
  out[2] = box_muller(r.z, r.w, false, sigma);
  out[3] = box_muller(r.z, r.w, true, sigma);
  if (D > 4) {
    u4 q = philox4x32_10(t_global, c, wn, 2u * iteration + 1u, seed_lo, seed_hi);
    out[D > 4 ? 4 : 0] = box_muller(q.x, q.y, false, sigma);
  }
}

// Box-Muller pair `pair` of one (sample, noise waypoint): pair 0 = joints (0,1), pair 1 =
// joints (2,3), pair 2 = joint 4 (second Philox call; *b unused).  Same values as noise_at.
DM_HD void noise_pair(uint32_t seed_lo, uint32_t seed_hi, float sigma, uint32_t iteration,
                      uint32_t c, uint32_t t_global, uint32_t wn, uint32_t pair, float *a,
                      float *b) {
  u4 r = philox4x32_10(t_global, c, wn, 2u * iteration + (pair >> 1), seed_lo, seed_hi);
  const uint32_t u = (pair & 1u) ? r.z : r.x, v = (pair & 1u) ? r.w : r.y;
  *a = box_muller(u, v, false, sigma);
  *b = box_muller(u, v, true, sigma);
}

} // namespace dm
