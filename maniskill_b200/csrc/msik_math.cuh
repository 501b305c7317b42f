// Small register-resident math for the EE-controller kernels (sm_100a).
// Everything here is per-thread fp32; no tensor cores (6x7 per-env systems are
// not a dense contraction -- see DESIGN.md).
#pragma once
#include <stdint.h>
#ifdef MSIK_HOST_TEST
// tests/ compile this header (and msik_ik.cuh, generated/chains.cuh) with g++ to run the per-env math of the
// kernels in fp32 on the CPU against the oracle: the CUDA qualifiers and intrinsics are mapped to libm.
#include <cmath>
#include <cstring>
#ifndef __device__
#define __device__
#endif
#ifndef __forceinline__
#define __forceinline__ inline
#endif
static inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
#else
#include <cuda_runtime.h>
#endif

namespace msik {

#ifdef MSIK_HOST_TEST
__device__ __forceinline__ float rcp_approx(float x) { return 1.0f / x; }
__device__ __forceinline__ float rsqrt_approx(float x) { return 1.0f / sqrtf(x); }
#else
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float rsqrt_approx(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
#endif
// 1/x to ~1 ulp: MUFU.RCP + one Newton step (2 FFMA)
__device__ __forceinline__ float rcp_nr(float x) {
  float r = rcp_approx(x);
  float e = fmaf(-x, r, 1.0f);
  return fmaf(r, e, r);
}
// sqrt(x) for x >= 0 to ~1 ulp without the libdevice slow path; sqrt(0) = 0
__device__ __forceinline__ float sqrt_nr(float x) {
  float r = rsqrt_approx(x + 1e-37f);  // FADD (fma pipe) rather than FMNMX (alu pipe); sqrt(0) = 0
  float s = x * r;                        // ~sqrt(x)
  float e = fmaf(-s, s, x);               // residual
  return fmaf(0.5f * r, e, s);
}

// sin/cos for joint angles (|x| < ~1e5 rad; joint limits keep it to a few radians).
// k = round(x / pi) by the magic-number trick (no F2I / FRND on the 4-lane xu pipe), two-constant
// Cody-Waite reduction r = x - k*pi in [-pi/2, pi/2], minimax polynomials on that interval
// (max abs error 1.3e-7 / 1.1e-7 in fp32), and sin(x) = (-1)^k sin(r), cos(x) = (-1)^k cos(r): the
// sign is one shift + two XORs, there is no quadrant select.  alu-pipe instructions cost two
// issue cycles on sm_100 (tools/microbench/ffma2.cu), so this form is ~35 % cheaper than the
// pi/2-reduction with selects.
__device__ __forceinline__ void sincos_joint(float x, float& s, float& c) {
#ifdef MSIK_USE_LIBDEVICE_SINCOS
  sincosf(x, &s, &c);
#elif defined(MSIK_USE_MUFU_SINCOS)
  s = __sinf(x); c = __cosf(x);  // experiment only: 2^-21.4 abs error, see profiles/README.md
#else
  const float MAGIC = 12582912.0f;  // 1.5 * 2^23
  float kf = fmaf(x, 0.318309886183790672f, MAGIC);
  const uint32_t sgn = __float_as_uint(kf) << 31;  // parity of k -> sign bit
  kf -= MAGIC;
  float r = fmaf(kf, -3.14159274101257324f, x);  // pi = hi + lo, hi = fp32(pi)
  r = fmaf(kf, 8.74227765734758577e-08f, r);
  const float r2 = r * r;
  float ps = fmaf(r2, 2.6000546e-06f, -1.9806615e-04f);
  ps = fmaf(ps, r2, 8.333017e-03f);
  ps = fmaf(ps, r2, -1.6666657e-01f);
  const float sr = fmaf(r * r2, ps, r);
  float pc = fmaf(r2, -2.6077106e-07f, 2.4761886e-05f);
  pc = fmaf(pc, r2, -1.3888404e-03f);
  pc = fmaf(pc, r2, 4.166664e-02f);
  pc = fmaf(pc, r2, -0.5f);
  const float cr = fmaf(pc, r2, 1.0f);
  s = __uint_as_float(__float_as_uint(sr) ^ sgn);
  c = __uint_as_float(__float_as_uint(cr) ^ sgn);
#endif
}

// The same function for TWO joint angles at once in packed fp32x2 arithmetic (sm_100 FFMA2 / FADD2 / FMUL2: one issue slot
// for two IEEE operations, each component rounded exactly like the scalar instruction -> bit-identical to two sincos_joint
// calls).  The kernels are bound by issue slots, not by the fma pipe (profiles/r02h_*: 73 % issue, ~50 % fma pipe), and the
// seven sincos of a 7-axis arm are a fifth of the loop: pairing them returns ~8 % of the instructions.
__device__ __forceinline__ void sincos_joint2(float x0, float x1, float& s0, float& c0, float& s1, float& c1) {
#if defined(MSIK_HOST_TEST) || defined(MSIK_NO_PACKED_SINCOS) || defined(MSIK_USE_LIBDEVICE_SINCOS) || defined(MSIK_USE_MUFU_SINCOS)
  sincos_joint(x0, s0, c0);
  sincos_joint(x1, s1, c1);
#else
#define MSIK_P2(v) make_float2(v, v)
  const float MAGIC = 12582912.0f;
  const float2 X = make_float2(x0, x1);
  float2 kf = __ffma2_rn(X, MSIK_P2(0.318309886183790672f), MSIK_P2(MAGIC));
  const uint32_t g0 = __float_as_uint(kf.x) << 31, g1 = __float_as_uint(kf.y) << 31;
  kf = __fadd2_rn(kf, MSIK_P2(-MAGIC));
  float2 r = __ffma2_rn(kf, MSIK_P2(-3.14159274101257324f), X);
  r = __ffma2_rn(kf, MSIK_P2(8.74227765734758577e-08f), r);
  const float2 r2 = __fmul2_rn(r, r);
  float2 ps = __ffma2_rn(r2, MSIK_P2(2.6000546e-06f), MSIK_P2(-1.9806615e-04f));
  ps = __ffma2_rn(ps, r2, MSIK_P2(8.333017e-03f));
  ps = __ffma2_rn(ps, r2, MSIK_P2(-1.6666657e-01f));
  const float2 sr = __ffma2_rn(__fmul2_rn(r, r2), ps, r);
  float2 pc = __ffma2_rn(r2, MSIK_P2(-2.6077106e-07f), MSIK_P2(2.4761886e-05f));
  pc = __ffma2_rn(pc, r2, MSIK_P2(-1.3888404e-03f));
  pc = __ffma2_rn(pc, r2, MSIK_P2(4.166664e-02f));
  pc = __ffma2_rn(pc, r2, MSIK_P2(-0.5f));
  const float2 cr = __ffma2_rn(pc, r2, MSIK_P2(1.0f));
#undef MSIK_P2
  s0 = __uint_as_float(__float_as_uint(sr.x) ^ g0);
  c0 = __uint_as_float(__float_as_uint(cr.x) ^ g0);
  s1 = __uint_as_float(__float_as_uint(sr.y) ^ g1);
  c1 = __uint_as_float(__float_as_uint(cr.y) ^ g1);
#endif
}

// (c0, c1) = (a0, a1) * b + (c0, c1): ONE packed FFMA2 on sm_100 (b broadcast), two scalar FFMAs elsewhere; each component
// is the same IEEE fma either way, so results do not depend on which form runs.  The kernels are issue-slot bound, the
// fma pipe half idle: every pair of same-shaped, independent FFMAs that shares one operand is worth packing.
// With three register operands an FFMA2 reads five registers, and register-file bandwidth is what these instructions queue
// on: measured, packing the Householder updates is 0.9 % SLOWER in the path-A kernel (21.32 vs 21.14 us per 2^20 envs) and
// 1.4 % faster in the iterative kernel (141.5 vs 143.5 us per 2^18 envs), so PACK is a per-call-site choice.
template <bool PACK>
__device__ __forceinline__ void fma2_b(float a0, float a1, float b, float& c0, float& c1) {
#if defined(MSIK_HOST_TEST) || defined(MSIK_NO_PACKED)
  c0 = fmaf(a0, b, c0);
  c1 = fmaf(a1, b, c1);
#else
  if constexpr (PACK) {
    const float2 v = __ffma2_rn(make_float2(a0, a1), make_float2(b, b), make_float2(c0, c1));
    c0 = v.x; c1 = v.y;
  } else {
    c0 = fmaf(a0, b, c0);
    c1 = fmaf(a1, b, c1);
  }
#endif
}

// sin / cos of every joint of a chain whose revolute joints are known at compile time (bit i of REV): revolute joints are
// taken two at a time through sincos_joint2, an odd one out through sincos_joint; other joints get (0, 1).
template <int N, uint32_t REV>
__device__ __forceinline__ void sincos_revolute_pairs(const float (&q)[N], float (&sn)[N], float (&cs)[N]) {
  int pending = -1;  // compile-time after unrolling
#pragma unroll
  for (int i = 0; i < N; ++i) {
    if ((REV >> i) & 1u) {
      if (pending < 0) {
        pending = i;
      } else {
        sincos_joint2(q[pending], q[i], sn[pending], cs[pending], sn[i], cs[i]);
        pending = -1;
      }
    } else {
      sn[i] = 0.f; cs[i] = 1.f;
    }
  }
  if (pending >= 0) sincos_joint(q[pending], sn[pending], cs[pending]);
}

struct Quat {
  float w, x, y, z;
};

// rotation_conversions.py:386-404
__device__ __forceinline__ Quat quat_raw_mul(const Quat& a, const Quat& b) {
  Quat o;
  o.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  o.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  o.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  o.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return o;
}
// rotation_conversions.py:371-383
__device__ __forceinline__ Quat quat_standardize(const Quat& q) {
  if (q.w < 0.f) return Quat{-q.w, -q.x, -q.y, -q.z};
  return q;
}
// rotation_conversions.py:407-421
__device__ __forceinline__ Quat quat_mul(const Quat& a, const Quat& b) {
  return quat_standardize(quat_raw_mul(a, b));
}
__device__ __forceinline__ Quat quat_conj(const Quat& q) { return Quat{q.w, -q.x, -q.y, -q.z}; }

// rotation_conversions.py:441-461: (q * (0,v)) * conj(q), imaginary part
__device__ __forceinline__ void quat_apply(const Quat& q, const float (&v)[3], float (&o)[3]) {
  Quat p{0.f, v[0], v[1], v[2]};
  Quat r = quat_raw_mul(quat_raw_mul(q, p), quat_conj(q));
  o[0] = r.x; o[1] = r.y; o[2] = r.z;
}

// rotation_conversions.py:105-164: candidate with the largest |component| (first
// maximum on ties, like torch.argmax), divided by 2*max(q_abs, 0.1), w >= 0.
__device__ __forceinline__ Quat mat_to_quat(const float (&m)[9]) {
  const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5], m20 = m[6],
              m21 = m[7], m22 = m[8];
  float a0 = 1.f + m00 + m11 + m22;
  float a1 = 1.f + m00 - m11 - m22;
  float a2 = 1.f - m00 + m11 - m22;
  float a3 = 1.f - m00 - m11 + m22;
  // argmax over sqrt(max(0,a_i)) == argmax over max(0,a_i)
  a0 = fmaxf(a0, 0.f); a1 = fmaxf(a1, 0.f); a2 = fmaxf(a2, 0.f); a3 = fmaxf(a3, 0.f);
  int idx = 0; float best = a0;
  if (a1 > best) { best = a1; idx = 1; }
  if (a2 > best) { best = a2; idx = 2; }
  if (a3 > best) { best = a3; idx = 3; }
  float qa = sqrt_nr(best);
  float inv = rcp_nr(2.f * fmaxf(qa, 0.1f));
  float sq = qa * qa;
  Quat q;
  if (idx == 0)      q = Quat{sq, m21 - m12, m02 - m20, m10 - m01};
  else if (idx == 1) q = Quat{m21 - m12, sq, m10 + m01, m02 + m20};
  else if (idx == 2) q = Quat{m02 - m20, m10 + m01, sq, m12 + m21};
  else               q = Quat{m10 - m01, m20 + m02, m21 + m12, sq};
  q.w *= inv; q.x *= inv; q.y *= inv; q.z *= inv;
  return quat_standardize(q);
}

// rotation_conversions.py:197-223, convention "XYZ": R = Rx(e0) Ry(e1) Rz(e2)
__device__ __forceinline__ void euler_xyz_to_mat(float e0, float e1, float e2, float (&m)[9]) {
  float s0, c0, s1, c1, s2, c2;
  sincosf(e0, &s0, &c0); sincosf(e1, &s1, &c1); sincosf(e2, &s2, &c2);
  // Rx*Ry = [[c1,0,s1],[s0 s1, c0, -s0 c1],[-c0 s1, s0, c0 c1]]
  m[0] = c1 * c2;               m[1] = -c1 * s2;              m[2] = s1;
  m[3] = s0 * s1 * c2 + c0 * s2; m[4] = -s0 * s1 * s2 + c0 * c2; m[5] = -s0 * c1;
  m[6] = -c0 * s1 * c2 + s0 * s2; m[7] = c0 * s1 * s2 + s0 * c2;  m[8] = c0 * c1;
}

// rotation_conversions.py:528-556
__device__ __forceinline__ void quat_to_axis_angle(const Quat& q, float (&o)[3]) {
  const float n2 = q.x * q.x + q.y * q.y + q.z * q.z;
  const float n = sqrt_nr(n2);
  const float half = atan2f(n, q.w);
  const float ang = 2.f * half;
  // sin(half) / ang with sin(atan2(n, w)) = n / sqrt(n^2 + w^2): no sinf (and none of its slow-path stack frame)
  float inv;  // 1 / (sin(half) / ang)
  if (fabsf(ang) < 1e-6f) {
    inv = rcp_nr(0.5f - ang * ang * (1.f / 48.f));
  } else {
    const float m2 = fmaf(q.w, q.w, n2);
    inv = ang * sqrt_nr(m2) * rcp_nr(n);
  }
  o[0] = q.x * inv; o[1] = q.y * inv; o[2] = q.z * inv;
}

// Minimum-norm solution of  A^T x = b  for A = J^T (NR joints x NC task dims, NR >= NC)
// by Householder QR of A:  x = Q [R^-T b; 0].  Equals pinv(J) b for full row rank J
// (kinematics.py:171) with error ~cond(J)*eps instead of the cond(J)^2*eps of the
// normal equations J J^T.  Pivots below rtol*max|pivot| are dropped (the analogue
// of torch.linalg.pinv's singular-value cut-off; only reached at exact singularities).
// SKIP: the LAST row of A is structurally zero in its first SKIP columns (a last joint whose
// tool offset lies on its own axis has a zero linear Jacobian column -- Panda, Fetch, xArm):
// the first SKIP reflectors then leave that row alone, which removes its FMAs at compile time.
template <int NR, int NC, int SKIP = 0>
__device__ __forceinline__ void min_norm_solve(float (&A)[NR][NC], const float (&b)[NC],
                                               float (&x)[NR], float rtol, float* pmin = nullptr) {
  static_assert(NR >= NC, "need at least as many controlled joints as task dimensions");
  static_assert(SKIP == 0 || NR > NC || SKIP < NC, "a zero last row needs NR > NC or a later column to pivot on");
  float tau[NC], rd[NC], rinv[NC];
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const int RE = (k < SKIP) ? NR - 1 : NR;  // rows k .. RE-1 take part in reflector k
    float sigma = 0.f;
#pragma unroll
    for (int r = k; r < RE; ++r) sigma = fmaf(A[r][k], A[r][k], sigma);
    const float akk = A[k][k];
    // MUFU.RSQ / MUFU.RCP are good to ~1 ulp: their error is a backward error of the reflector of the same
    // size as the fp32 roundings around it, so no Newton step (measured: same parity, 3 % faster)
    const float nrm = sigma * rsqrt_approx(sigma + 1e-37f);  // sqrt(0) = 0
    // alpha = -copysign(nrm, akk): one LOP3 (nrm >= 0)
    const float alpha = __uint_as_float(__float_as_uint(nrm) | (~__float_as_uint(akk) & 0x80000000u));
    const float v0 = akk - alpha;
    const float den = -alpha * v0;  // = sigma + |akk| * nrm >= 0
    // a zero column has v = 0: its dot products vanish, so t only has to stay finite
    const float t = rcp_approx(den + 1e-37f);
    tau[k] = t;
    rd[k] = alpha;
    rinv[k] = -v0 * t;  // 1 / alpha without a second reciprocal (t = -1 / (alpha v0)); 0 for a zero column
    A[k][k] = v0;
#pragma unroll
    for (int j = k + 1; j < NC; ++j) {
      float s = 0.f;
#pragma unroll
      for (int r = k; r < RE; ++r) s = fmaf(A[r][k], A[r][j], s);
      s *= t;
#pragma unroll
      for (int r = k; r < RE; ++r) A[r][j] = fmaf(-s, A[r][k], A[r][j]);
    }
  }
  float rmax = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) rmax = fmaxf(rmax, fabsf(rd[k]));
  const float thr = rtol * rmax;
  if (pmin) {  // smallest |pivot|: an upper bound of sigma_min(J), the near-singularity guard of the iterative solver
    float m = fabsf(rd[0]);
#pragma unroll
    for (int k = 1; k < NC; ++k) m = fminf(m, fabsf(rd[k]));
    *pmin = m;
  }
  float y[NC];
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    float acc = b[k];
#pragma unroll
    for (int i = 0; i < k; ++i) acc = fmaf(-A[i][k], y[i], acc);
    y[k] = fabsf(rd[k]) > thr ? acc * rinv[k] : 0.f;
  }
  // x = H_0 ... H_{NC-1} [y; 0]
#pragma unroll
  for (int r = 0; r < NR; ++r) x[r] = r < NC ? y[r] : 0.f;
#pragma unroll
  for (int k = NC - 1; k >= 0; --k) {
    const int RE = (k < SKIP) ? NR - 1 : NR;
    float s = 0.f;
    // entries below row NC-1 are still zero when the first reflector is applied
    const int rend = (k == NC - 1) ? (NC < RE ? NC : RE) : RE;
#pragma unroll
    for (int r = k; r < rend; ++r) s = fmaf(A[r][k], x[r], s);
    s *= tau[k];
#pragma unroll
    for (int r = k; r < RE; ++r) x[r] = fmaf(-s, A[r][k], x[r]);
  }
}

// Reduced path-A solve (generated chains, see maniskill_b200/chaingen.py gen_reduced):
//     minimise |u|^2 + |v|^2   subject to   A u = a1 (4 rows),   B u + v = a2 (2 rows)
// with At = A^T (NU x 4), NU = controlled joints - 2.  Householder QR of At gives the particular solution
// u_p = Q [R^-T a1; 0] and, for NU = 5, the unit null vector n = Q e_5 of A (u_p is orthogonal to n); then
// u = u_p + n t with t = (Bn)^T (a2 - B u_p) / (1 + |Bn|^2) and v = a2 - B u.  Same solution as pinv(J) a
// (the row operations that produced A, B do not change the solution set of J dq = a), a quarter of the
// reflector work of the dense NR x 6 factorisation.  Pivots below rtol * max|pivot| are dropped as in
// min_norm_solve.
// SKIP: the LAST row of At is structurally zero in its first SKIP columns (the generator orders unknowns and task
// rows that way), so the first SKIP reflectors leave it alone; ZB bit r * NU + i: Bm[r][i] is structurally zero.
template <int NU, int SKIP = 0, uint32_t ZB = 0u, bool PACK = false>
__device__ __forceinline__ void reduced_min_norm(float (&At)[NU][4], const float (&Bm)[2][NU], const float (&a1)[4],
                                                 const float (&a2)[2], float (&u)[NU], float (&v)[2], float rtol,
                                                 float* pmin = nullptr) {
  static_assert(NU == 4 || NU == 5, "reduced solve: 6 or 7 controlled joints");
  static_assert(SKIP >= 0 && SKIP < 4, "a row that is zero in every column makes A rank deficient");
  constexpr int NC = 4;
  float tau[NC], rd[NC], rinv[NC];
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const int RE = (k < SKIP) ? NU - 1 : NU;  // rows k .. RE-1 take part in reflector k
    float sigma = 0.f;
#pragma unroll
    for (int r = k; r < RE; ++r) sigma = fmaf(At[r][k], At[r][k], sigma);
    const float akk = At[k][k];
    const float nrm = sigma * rsqrt_approx(sigma + 1e-37f);
    const float alpha = __uint_as_float(__float_as_uint(nrm) | (~__float_as_uint(akk) & 0x80000000u));
    const float v0 = akk - alpha;
    const float den = -alpha * v0;
    const float t = rcp_approx(den + 1e-37f);
    tau[k] = t;
    rd[k] = alpha;
    rinv[k] = -v0 * t;
    At[k][k] = v0;
    // columns k+1 .. 3 of the trailing block: columns 2 and 3 go as a PAIR (fma2_b: same per-column operation order as the
    // scalar loop), column 1 (k = 0) and column 3 (k = 2) alone
#pragma unroll
    for (int j = k + 1; j < NC; ++j) {
      if (k <= 1 && j >= 2) continue;  // handled as the pair below
      float s = 0.f;
#pragma unroll
      for (int r = k; r < RE; ++r) s = fmaf(At[r][k], At[r][j], s);
      s *= t;
#pragma unroll
      for (int r = k; r < RE; ++r) At[r][j] = fmaf(-s, At[r][k], At[r][j]);
    }
    if (k <= 1) {
      float s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int r = k; r < RE; ++r) fma2_b<PACK>(At[r][2], At[r][3], At[r][k], s2, s3);
      s2 *= t; s3 *= t;
#pragma unroll
      for (int r = k; r < RE; ++r) fma2_b<PACK>(-s2, -s3, At[r][k], At[r][2], At[r][3]);
    }
  }
  float rmax = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) rmax = fmaxf(rmax, fabsf(rd[k]));
  const float thr = rtol * rmax;
  // rank cut-off: a pivot below thr is dropped (its y component is zero).  Only exact singularities get there, so the
  // four compare + select pairs (alu pipe: two issue cycles each on sm_100) sit behind ONE test of the smallest pivot
  float y[NC];
  const float pm = fminf(fminf(fabsf(rd[0]), fabsf(rd[1])), fminf(fabsf(rd[2]), fabsf(rd[3])));
  if (pm > thr) {
#pragma unroll
    for (int k = 0; k < NC; ++k) {
      float acc = a1[k];
#pragma unroll
      for (int i = 0; i < k; ++i) acc = fmaf(-At[i][k], y[i], acc);
      y[k] = acc * rinv[k];
    }
  } else {
#pragma unroll
    for (int k = 0; k < NC; ++k) {
      float acc = a1[k];
#pragma unroll
      for (int i = 0; i < k; ++i) acc = fmaf(-At[i][k], y[i], acc);
      y[k] = fabsf(rd[k]) > thr ? acc * rinv[k] : 0.f;
    }
  }
  if (pmin) *pmin = pm;
#pragma unroll
  for (int r = 0; r < NU; ++r) u[r] = r < NC ? y[r] : 0.f;
  float n[NU];
#pragma unroll
  for (int r = 0; r < NU; ++r) n[r] = r == NU - 1 ? 1.f : 0.f;
#pragma unroll
  for (int k = NC - 1; k >= 0; --k) {
    const int RE = (k < SKIP) ? NU - 1 : NU;
    const int rend = (k == NC - 1) ? NC : RE;  // u[NC..] is still zero under the first reflector applied
    if (NU > NC && k < NC - 1) {
      // u and the null vector n go through the reflector as a PAIR: n = H_0 .. H_3 e_NU, n[k] is still zero here
      float s = At[k][k] * u[k], sn_ = 0.f;  // (= fmaf(At[k][k], u[k], 0): the r = k term of the scalar loop)
#pragma unroll
      for (int r = k + 1; r < RE; ++r) fma2_b<PACK>(u[r], n[r], At[r][k], s, sn_);
      s *= tau[k]; sn_ *= tau[k];
#pragma unroll
      for (int r = k; r < RE; ++r) fma2_b<PACK>(-s, -sn_, At[r][k], u[r], n[r]);
    } else {
      float s = 0.f;
#pragma unroll
      for (int r = k; r < rend; ++r) s = fmaf(At[r][k], u[r], s);
      s *= tau[k];
#pragma unroll
      for (int r = k; r < RE; ++r) u[r] = fmaf(-s, At[r][k], u[r]);
      if (NU > NC) {  // k == NC - 1: under this reflector only n[NU-1] = 1 is non-zero
        const float sn_ = At[NU - 1][k] * tau[k];
#pragma unroll
        for (int r = k; r < RE; ++r) n[r] = fmaf(-sn_, At[r][k], n[r]);
      }
    }
  }
  // w = a2 - B u_p
  float w0 = a2[0], w1 = a2[1];
#pragma unroll
  for (int r = 0; r < NU; ++r) {
    const bool z0 = (ZB >> r) & 1u, z1 = (ZB >> (NU + r)) & 1u;
    if (!z0 && !z1) fma2_b<PACK>(-Bm[0][r], -Bm[1][r], u[r], w0, w1);
    else if (!z0) w0 = fmaf(-Bm[0][r], u[r], w0);
    else if (!z1) w1 = fmaf(-Bm[1][r], u[r], w1);
  }
  if (NU > NC) {
    float g0 = 0.f, g1 = 0.f;  // B n
#pragma unroll
    for (int r = 0; r < NU; ++r) {
      const bool z0 = (ZB >> r) & 1u, z1 = (ZB >> (NU + r)) & 1u;
      if (!z0 && !z1) fma2_b<PACK>(Bm[0][r], Bm[1][r], n[r], g0, g1);
      else if (!z0) g0 = fmaf(Bm[0][r], n[r], g0);
      else if (!z1) g1 = fmaf(Bm[1][r], n[r], g1);
    }
    const float num = fmaf(g0, w0, g1 * w1);
    const float den = fmaf(g0, g0, fmaf(g1, g1, 1.f));
    const float t = num * rcp_nr(den);
#pragma unroll
    for (int r = 0; r < NU; ++r) u[r] = fmaf(n[r], t, u[r]);
    w0 = fmaf(-g0, t, w0); w1 = fmaf(-g1, t, w1);
  }
  v[0] = w0; v[1] = w1;
}

// Damped least squares  x = A (A^T A + lambda^2 I)^-1 b  for A = J^T (NR joints x NC task dims), i.e. the update
// J^T (J J^T + lambda^2 I)^-1 dx that pk.PseudoInverseIK builds (kinematics.py:110-117), WITHOUT forming
// J J^T: Householder QR of the augmented matrix [A; lambda I] = Q~ [R~; 0] gives R~^T R~ = A^T A + lambda^2 I, hence
//     x = [I 0] Q~ [R~^-T b; 0],
// with error ~ cond(R~) eps instead of the cond^2 of the normal equations.  Reflector k spans the rows k.. of A plus the
// rows 0..k of the lambda block; the fill-in of that block is the upper triangle L[i][j], i <= j.  lambda > 0 makes
// every pivot non-zero, so there is no rank cut-off here: near a singularity the damping does what it is there for.
template <int NR, int NC>
__device__ __forceinline__ void dls_solve(float (&A)[NR][NC], const float (&b)[NC], float (&x)[NR], float lambda) {
  float L[NC][NC];  // only i <= j is ever read
  float tau[NC], rinv[NC];
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    L[k][k] = lambda;
    float sigma = lambda * lambda;
#pragma unroll
    for (int r = k; r < NR; ++r) sigma = fmaf(A[r][k], A[r][k], sigma);
#pragma unroll
    for (int i = 0; i < k; ++i) sigma = fmaf(L[i][k], L[i][k], sigma);
    const float akk = A[k][k];
    const float nrm = sigma * rsqrt_approx(sigma);
    const float alpha = __uint_as_float(__float_as_uint(nrm) | (~__float_as_uint(akk) & 0x80000000u));  // -copysign(nrm, akk)
    const float v0 = akk - alpha;
    const float t = rcp_approx(-alpha * v0);  // = 1 / (sigma + |akk| nrm) > 0
    tau[k] = t;
    rinv[k] = -v0 * t;  // 1 / alpha
    A[k][k] = v0;
#pragma unroll
    for (int j = k + 1; j < NC; ++j) {
      float s = 0.f;
#pragma unroll
      for (int r = k; r < NR; ++r) s = fmaf(A[r][k], A[r][j], s);
#pragma unroll
      for (int i = 0; i < k; ++i) s = fmaf(L[i][k], L[i][j], s);  // row k of the block is still zero in column j
      s *= t;
#pragma unroll
      for (int r = k; r < NR; ++r) A[r][j] = fmaf(-s, A[r][k], A[r][j]);
#pragma unroll
      for (int i = 0; i < k; ++i) L[i][j] = fmaf(-s, L[i][k], L[i][j]);
      L[k][j] = -s * lambda;
    }
  }
  float w[NC];  // R~^T w = b
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    float acc = b[k];
#pragma unroll
    for (int i = 0; i < k; ++i) acc = fmaf(-A[i][k], w[i], acc);
    w[k] = acc * rinv[k];
  }
  float xb[NC];  // lambda-block part of Q~ [w; 0]; needed by the reflectors still to come, dropped at the end
#pragma unroll
  for (int r = 0; r < NR; ++r) x[r] = r < NC ? w[r] : 0.f;
#pragma unroll
  for (int i = 0; i < NC; ++i) xb[i] = 0.f;
#pragma unroll
  for (int k = NC - 1; k >= 0; --k) {
    float s = 0.f;
    const int rend = (k == NC - 1) ? NC : NR;  // x[NC..] and xb[] are still zero under the first reflector applied
#pragma unroll
    for (int r = k; r < rend; ++r) s = fmaf(A[r][k], x[r], s);
    if (k < NC - 1) {
#pragma unroll
      for (int i = 0; i <= k; ++i) s = fmaf(L[i][k], xb[i], s);
    }
    s *= tau[k];
#pragma unroll
    for (int r = k; r < NR; ++r) x[r] = fmaf(-s, A[r][k], x[r]);
#pragma unroll
    for (int i = 0; i < k; ++i) xb[i] = fmaf(-s, L[i][k], xb[i]);  // xb[k] is never read again
  }
}

// Rotation vector of E = T C^T (the rotation that takes orientation C to orientation T; row-major rotation matrices in
// one common frame) = axis_angle(q_T (x) conj(q_C)), the orientation error of the iterative solver.
// Small angles (<= 45 deg, every iteration after the first few): E is never formed -- 2 sin(th) a = (E21 - E12, ...)
// and tr E come straight from the rows of T and C, th / sin(th) = 2 h(u) / (1 + cos th) with u = tan^2(th / 2) and
// h(u) = atan(sqrt u) / sqrt u as a degree-5 polynomial (3.5e-8 in fp32 Horner): no square root, no atan2f, no
// quaternion.  Larger angles: E -> quaternion -> axis-angle exactly as the reference's helpers do it
// (rotation_conversions.py:105-164,528-556).
__device__ __forceinline__ void rot_err(const float (&T)[9], const float (&C)[9], float (&w)[3]) {
  const float ax = fmaf(T[6], C[3], fmaf(T[7], C[4], T[8] * C[5])) - fmaf(T[3], C[6], fmaf(T[4], C[7], T[5] * C[8]));
  const float ay = fmaf(T[0], C[6], fmaf(T[1], C[7], T[2] * C[8])) - fmaf(T[6], C[0], fmaf(T[7], C[1], T[8] * C[2]));
  const float az = fmaf(T[3], C[0], fmaf(T[4], C[1], T[5] * C[2])) - fmaf(T[0], C[3], fmaf(T[1], C[4], T[2] * C[5]));
  float tr = T[0] * C[0];
#pragma unroll
  for (int k = 1; k < 9; ++k) tr = fmaf(T[k], C[k], tr);
  const float d = 1.f + tr;  // 2 (1 + cos th)
  if (d >= 3.4142137f) {     // th <= 45 deg
    const float r = rcp_nr(d);
    const float u = fmaf(ax, ax, fmaf(ay, ay, az * az)) * (r * r);  // tan^2(th / 2)
    float h = fmaf(u, -0.060263053f, 0.10569829f);
    h = fmaf(h, u, -0.14239533f);
    h = fmaf(h, u, 0.19998182f);
    h = fmaf(h, u, -0.33333308f);
    h = fmaf(h, u, 1.0f);
    const float g = 2.f * h * r;
    w[0] = ax * g; w[1] = ay * g; w[2] = az * g;
    return;
  }
  float E[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) E[3 * i + j] = fmaf(T[3 * i], C[3 * j], fmaf(T[3 * i + 1], C[3 * j + 1], T[3 * i + 2] * C[3 * j + 2]));
  quat_to_axis_angle(mat_to_quat(E), w);
}

// unit quaternion (w, x, y, z) -> row-major rotation matrix (rotation_conversions.py:41-72 for a unit input)
__device__ __forceinline__ void quat_to_mat(const Quat& q, float (&m)[9]) {
  const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z, xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
  const float wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
  m[0] = 1.f - 2.f * (yy + zz); m[1] = 2.f * (xy - wz);       m[2] = 2.f * (xz + wy);
  m[3] = 2.f * (xy + wz);       m[4] = 1.f - 2.f * (xx + zz); m[5] = 2.f * (yz - wx);
  m[6] = 2.f * (xz - wy);       m[7] = 2.f * (yz + wx);       m[8] = 1.f - 2.f * (xx + yy);
}

}  // namespace msik
