/*
 * ik_oracle.c -- fp64 C restatement of the end-effector controller hot path.
 *
 * TEST INFRASTRUCTURE / TIMED CPU BASELINE ONLY.  Linked or executed only from
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  The product (maniskill_b200/) never touches it.
 *
 * PARITY UNPINNED at the third-party boundary: the arithmetic of this path lives in
 * pytorch_kinematics==0.7.5 (reference setup.py:46) and SAPIEN's PinocchioModel
 * (sapien>=3.0.0.b1, setup.py:50-51), both absent from /root/reference and not
 * installable.  Restated from their published algorithms and the reference's own
 * call sites; checked against oracle/oracle.py (numpy) and the known-answer
 * vectors of SURVEY.md Appendix A.1 in tests/test_oracle.py.
 *
 *   orc_fk_jac        SURVEY Appendix A.1 (pk chain.jacobian, kinematics.py:157)
 *   orc_delta_step    q0 + pinv(J) a, SVD cut-off max(m,n)*eps_fp32 (kinematics.py:156-172)
 *   orc_iterative     pk.PseudoInverseIK.solve (kinematics.py:110-117,144-155)
 *   orc_clik          PinocchioModel.compute_inverse_kinematics (kinematics.py:173-186)
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define ORC_MAXJ 16
#define ORC_FIXED 0
#define ORC_REVOLUTE 1
#define ORC_PRISMATIC 2

typedef struct orc_frame {
  int32_t type;
  double xyz[3];
  double rpy[3];
  double axis[3];
} orc_frame;

static void mat3_mul(const double* A, const double* B, double* C) {
  double T[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) T[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(C, T, sizeof T);
}
static void mat3_vec(const double* A, const double* v, double* o) {
  double t[3];
  for (int i = 0; i < 3; ++i) t[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
  memcpy(o, t, sizeof t);
}
static void mat3_tvec(const double* A, const double* v, double* o) { /* A^T v */
  double t[3];
  for (int i = 0; i < 3; ++i) t[i] = A[i] * v[0] + A[3 + i] * v[1] + A[6 + i] * v[2];
  memcpy(o, t, sizeof t);
}
static void cross3(const double* a, const double* b, double* o) {
  double t[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
  memcpy(o, t, sizeof t);
}
static void rpy_mat(const double* rpy, double* R) {
  const double cr = cos(rpy[0]), sr = sin(rpy[0]), cp = cos(rpy[1]), sp = sin(rpy[1]), cy = cos(rpy[2]), sy = sin(rpy[2]);
  R[0] = cy * cp; R[1] = cy * sp * sr - sy * cr; R[2] = cy * sp * cr + sy * sr;
  R[3] = sy * cp; R[4] = sy * sp * sr + cy * cr; R[5] = sy * sp * cr - cy * sr;
  R[6] = -sp;     R[7] = cp * sr;                R[8] = cp * cr;
}
static void axis_rot(const double* a, double th, double* R) {
  const double c = cos(th), s = sin(th), C = 1 - c, x = a[0], y = a[1], z = a[2];
  R[0] = c + x * x * C;     R[1] = x * y * C - z * s; R[2] = x * z * C + y * s;
  R[3] = y * x * C + z * s; R[4] = c + y * y * C;     R[5] = y * z * C - x * s;
  R[6] = z * x * C - y * s; R[7] = z * y * C + x * s; R[8] = c + z * z * C;
}

/* FK + base-frame geometric Jacobian at the EE origin; J is 6 x n row-major, rows [lin; ang]. */
int orc_fk_jac(const orc_frame* fr, int nf, const double* q, double* J, double* R_out, double* p_out) {
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, p[3] = {0, 0, 0};
  double ax[ORC_MAXJ][3], org[ORC_MAXJ][3];
  int ty[ORC_MAXJ];
  int k = 0;
  for (int f = 0; f < nf; ++f) {
    double Ro[9], t[3];
    rpy_mat(fr[f].rpy, Ro);
    mat3_vec(R, fr[f].xyz, t);
    p[0] += t[0]; p[1] += t[1]; p[2] += t[2];
    mat3_mul(R, Ro, R);
    if (fr[f].type == ORC_FIXED) continue;
    if (k >= ORC_MAXJ) return -1;
    double a[3];
    const double na = sqrt(fr[f].axis[0] * fr[f].axis[0] + fr[f].axis[1] * fr[f].axis[1] + fr[f].axis[2] * fr[f].axis[2]);
    for (int i = 0; i < 3; ++i) a[i] = fr[f].axis[i] / na;
    mat3_vec(R, a, ax[k]);
    memcpy(org[k], p, sizeof p);
    ty[k] = fr[f].type;
    if (fr[f].type == ORC_REVOLUTE) {
      double Rq[9];
      axis_rot(a, q[k], Rq);
      mat3_mul(R, Rq, R);
    } else {
      for (int i = 0; i < 3; ++i) p[i] += ax[k][i] * q[k];
    }
    ++k;
  }
  const int n = k;
  if (J) {
    for (int i = 0; i < n; ++i) {
      if (ty[i] == ORC_REVOLUTE) {
        double d[3] = {p[0] - org[i][0], p[1] - org[i][1], p[2] - org[i][2]}, l[3];
        cross3(ax[i], d, l);
        for (int r = 0; r < 3; ++r) { J[r * n + i] = l[r]; J[(3 + r) * n + i] = ax[i][r]; }
      } else {
        for (int r = 0; r < 3; ++r) { J[r * n + i] = ax[i][r]; J[(3 + r) * n + i] = 0.0; }
      }
    }
  }
  if (R_out) memcpy(R_out, R, sizeof R);
  if (p_out) memcpy(p_out, p, sizeof p);
  return n;
}

/* x = pinv(J) b for J (m x n, m <= 6) by one-sided Jacobi SVD of A = J^T (n x m);
 * singular values <= rtol * sigma_max are dropped (torch.linalg.pinv semantics). */
static void pinv_apply(const double* J, int m, int n, const double* b, double rtol, double* x) {
  double A[ORC_MAXJ][6], V[6][6];
  for (int r = 0; r < n; ++r) for (int c = 0; c < m; ++c) A[r][c] = J[c * n + r];
  for (int i = 0; i < m; ++i) for (int j = 0; j < m; ++j) V[i][j] = i == j;
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0;
    for (int i = 0; i < m - 1; ++i)
      for (int j = i + 1; j < m; ++j) {
        double al = 0, be = 0, ga = 0;
        for (int r = 0; r < n; ++r) { al += A[r][i] * A[r][i]; be += A[r][j] * A[r][j]; ga += A[r][i] * A[r][j]; }
        if (fabs(ga) <= 1e-300 || fabs(ga) <= 1e-17 * sqrt(al * be)) continue;
        off = fmax(off, fabs(ga) / sqrt(al * be));
        const double zeta = (be - al) / (2 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1 + zeta * zeta));
        const double c = 1 / sqrt(1 + t * t), s = c * t;
        for (int r = 0; r < n; ++r) { const double u = A[r][i], v = A[r][j]; A[r][i] = c * u - s * v; A[r][j] = s * u + c * v; }
        for (int r = 0; r < m; ++r) { const double u = V[r][i], v = V[r][j]; V[r][i] = c * u - s * v; V[r][j] = s * u + c * v; }
      }
    if (off < 1e-15) break;
  }
  double s2[6], smax2 = 0;
  for (int i = 0; i < m; ++i) { s2[i] = 0; for (int r = 0; r < n; ++r) s2[i] += A[r][i] * A[r][i]; smax2 = fmax(smax2, s2[i]); }
  for (int r = 0; r < n; ++r) x[r] = 0;
  for (int i = 0; i < m; ++i) {
    if (s2[i] <= rtol * rtol * smax2) continue;
    double vb = 0;
    for (int c = 0; c < m; ++c) vb += V[c][i] * b[c];
    const double w = vb / s2[i];
    for (int r = 0; r < n; ++r) x[r] += A[r][i] * w;
  }
}

/* SPD solve (6x6) by Cholesky */
static int chol_solve6(double* A, const double* b, double* x) {
  double Lm[6][6] = {{0}};
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = A[6 * i + j];
      for (int k = 0; k < j; ++k) s -= Lm[i][k] * Lm[j][k];
      if (i == j) { if (s <= 0) return -1; Lm[i][i] = sqrt(s); } else Lm[i][j] = s / Lm[j][j];
    }
  double y[6];
  for (int i = 0; i < 6; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= Lm[i][k] * y[k]; y[i] = s / Lm[i][i]; }
  for (int i = 5; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < 6; ++k) s -= Lm[k][i] * x[k]; x[i] = s / Lm[i][i]; }
  return 0;
}

/* path A.  mask: n entries (1 = controlled); out: one value per controlled joint. */
int orc_delta_step(const orc_frame* fr, int nf, const double* q, const double* a, int nc, const uint8_t* mask, double* out) {
  double J[6 * ORC_MAXJ];
  const int n = orc_fk_jac(fr, nf, q, J, 0, 0);
  if (n < 0) return -1;
  double Jm[6 * ORC_MAXJ];
  int c = 0, cols[ORC_MAXJ];
  for (int i = 0; i < n; ++i) if (!mask || mask[i]) cols[c++] = i;
  for (int r = 0; r < nc; ++r) for (int j = 0; j < c; ++j) Jm[r * c + j] = J[r * n + cols[j]];
  double dq[ORC_MAXJ];
  const double rtol = (double)(nc > c ? nc : c) * 1.1920928955078125e-07;
  pinv_apply(Jm, nc, c, a, rtol, dq);
  for (int j = 0; j < c; ++j) out[j] = q[cols[j]] + dq[j];
  return c;
}

static void mat_to_quat(const double* m, double* q) { /* rotation_conversions.py:105-164 */
  const double a[4] = {1 + m[0] + m[4] + m[8], 1 + m[0] - m[4] - m[8], 1 - m[0] + m[4] - m[8], 1 - m[0] - m[4] + m[8]};
  double qa[4];
  int idx = 0;
  for (int i = 0; i < 4; ++i) { qa[i] = a[i] > 0 ? sqrt(a[i]) : 0; if (qa[i] > qa[idx]) idx = i; }
  double c[4];
  switch (idx) {
    case 0: c[0] = qa[0] * qa[0]; c[1] = m[7] - m[5]; c[2] = m[2] - m[6]; c[3] = m[3] - m[1]; break;
    case 1: c[0] = m[7] - m[5]; c[1] = qa[1] * qa[1]; c[2] = m[3] + m[1]; c[3] = m[2] + m[6]; break;
    case 2: c[0] = m[2] - m[6]; c[1] = m[3] + m[1]; c[2] = qa[2] * qa[2]; c[3] = m[5] + m[7]; break;
    default: c[0] = m[3] - m[1]; c[1] = m[6] + m[2]; c[2] = m[7] + m[5]; c[3] = qa[3] * qa[3]; break;
  }
  const double d = 2.0 * fmax(qa[idx], 0.1);
  const double sg = c[0] < 0 ? -1.0 : 1.0;
  for (int i = 0; i < 4; ++i) q[i] = sg * c[i] / d;
}
static void quat_mul_std(const double* a, const double* b, double* o) { /* :386-421 */
  double t[4] = {a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                 a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]};
  const double sg = t[0] < 0 ? -1.0 : 1.0;
  for (int i = 0; i < 4; ++i) o[i] = sg * t[i];
}
static void quat_axis_angle(const double* q, double* o) { /* :528-556 */
  const double n = sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double half = atan2(n, q[0]), ang = 2 * half;
  const double k = fabs(ang) < 1e-6 ? 0.5 - ang * ang / 48 : sin(half) / ang;
  for (int i = 0; i < 3; ++i) o[i] = q[1 + i] / k;
}

/* path B.  tgt = [p, q wxyz].  Returns iterations used; out per controlled joint. */
int orc_iterative(const orc_frame* fr, int nf, const double* q0, const double* tgt, const uint8_t* mask, int max_it,
                  double lr, double reg, double pos_tol, double rot_tol, int patience, double* out) {
  double q[ORC_MAXJ], sol[ORC_MAXJ], J[6 * ORC_MAXJ], R[9], p[3];
  int n = orc_fk_jac(fr, nf, q0, 0, 0, 0);
  if (n < 0) return -1;
  memcpy(q, q0, sizeof(double) * n);
  memcpy(sol, q0, sizeof(double) * n);
  double tq[4];
  { double nn = sqrt(tgt[3] * tgt[3] + tgt[4] * tgt[4] + tgt[5] * tgt[5] + tgt[6] * tgt[6]); double sg = tgt[3] < 0 ? -1 : 1;
    for (int i = 0; i < 4; ++i) tq[i] = sg * tgt[3 + i] / nn; }
  double err_min = INFINITY;
  int noimp = 0, it = 0;
  for (; it < max_it; ++it) {
    orc_fk_jac(fr, nf, q, J, R, p);
    for (int i = 0; i < n; ++i) if (mask && !mask[i]) for (int r = 0; r < 6; ++r) J[r * n + i] = 0;
    double dx[6], cq[4], ci[4], dqt[4];
    for (int i = 0; i < 3; ++i) dx[i] = tgt[i] - p[i];
    mat_to_quat(R, cq);
    ci[0] = cq[0]; ci[1] = -cq[1]; ci[2] = -cq[2]; ci[3] = -cq[3];
    quat_mul_std(tq, ci, dqt);
    quat_axis_angle(dqt, dx + 3);
    double A[36], y[6];
    for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) { double s = 0; for (int i = 0; i < n; ++i) s += J[r * n + i] * J[c * n + i]; A[6 * r + c] = s + (r == c ? reg : 0); }
    if (chol_solve6(A, dx, y)) break;
    for (int i = 0; i < n; ++i) { double s = 0; for (int r = 0; r < 6; ++r) s += J[r * n + i] * y[r]; q[i] += lr * s; sol[i] = q[i]; }
    const double ep = sqrt(dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2]), er = sqrt(dx[3] * dx[3] + dx[4] * dx[4] + dx[5] * dx[5]);
    const double err = sqrt(ep * ep + er * er);
    if (err < err_min) { err_min = err; noimp = 0; } else ++noimp;
    if (ep < pos_tol && er < rot_tol) { ++it; break; }
    if (patience >= 0 && noimp > patience) { ++it; break; }
  }
  int c = 0;
  for (int i = 0; i < n; ++i) if (!mask || mask[i]) out[c++] = sol[i];
  return it;
}

static void log3(const double* R, double* w) {
  double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);
  c = c > 1 ? 1 : (c < -1 ? -1 : c);
  const double th = acos(c);
  double v[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
  if (th < 1e-8) { for (int i = 0; i < 3; ++i) w[i] = 0.5 * v[i]; return; }
  if (th > M_PI - 1e-6) {
    double A[9]; for (int i = 0; i < 9; ++i) A[i] = 0.5 * (R[i] + (i % 4 == 0));
    int k = 0; for (int i = 1; i < 3; ++i) if (A[4 * i] > A[4 * k]) k = i;
    double ax[3]; const double d = sqrt(fmax(A[4 * k], 1e-300));
    for (int i = 0; i < 3; ++i) ax[i] = A[3 * i + k] / d;
    if (v[0] * ax[0] + v[1] * ax[1] + v[2] * ax[2] < 0) for (int i = 0; i < 3; ++i) ax[i] = -ax[i];
    for (int i = 0; i < 3; ++i) w[i] = ax[i] * th;
    return;
  }
  const double k = th / (2 * sin(th));
  for (int i = 0; i < 3; ++i) w[i] = v[i] * k;
}
static void log6(const double* R, const double* p, double* out) { /* Motion order [v; w] */
  double w[3], wp[3], wwp[3];
  log3(R, w);
  const double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  const double beta = th < 1e-6 ? 1.0 / 12 + th * th / 720 : (1 - th * sin(th) / (2 * (1 - cos(th)))) / (th * th);
  cross3(w, p, wp);
  cross3(w, wp, wwp);
  for (int i = 0; i < 3; ++i) { out[i] = p[i] - 0.5 * wp[i] + beta * wwp[i]; out[3 + i] = w[i]; }
}

/* path C: CLIK.  Returns 1 on success, 0 on failure; out = best q (all n chain joints). */
int orc_clik(const orc_frame* fr, int nf, const double* q0, const double* tgt, const uint8_t* mask, int max_it, double eps,
             double dt, double damp, double* out, double* best_err_norm) {
  double q[ORC_MAXJ], J[6 * ORC_MAXJ], R[9], p[3], Rt[9];
  int n = orc_fk_jac(fr, nf, q0, 0, 0, 0);
  if (n < 0) return -1;
  memcpy(q, q0, sizeof(double) * n);
  memcpy(out, q0, sizeof(double) * n);
  { const double nn = sqrt(tgt[3] * tgt[3] + tgt[4] * tgt[4] + tgt[5] * tgt[5] + tgt[6] * tgt[6]);
    const double r = tgt[3] / nn, i = tgt[4] / nn, j = tgt[5] / nn, k = tgt[6] / nn;
    Rt[0] = 1 - 2 * (j * j + k * k); Rt[1] = 2 * (i * j - k * r); Rt[2] = 2 * (i * k + j * r);
    Rt[3] = 2 * (i * j + k * r); Rt[4] = 1 - 2 * (i * i + k * k); Rt[5] = 2 * (j * k - i * r);
    Rt[6] = 2 * (i * k - j * r); Rt[7] = 2 * (j * k + i * r); Rt[8] = 1 - 2 * (i * i + j * j); }
  double min_err = 1e10;
  int success = 0;
  for (int it = 0;; ++it) {
    orc_fk_jac(fr, nf, q, J, R, p);
    double Rd[9], RtT[9], d[3], pd[3], err[6];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) RtT[3 * i + j] = Rt[3 * j + i];
    mat3_mul(RtT, R, Rd);
    for (int i = 0; i < 3; ++i) d[i] = p[i] - tgt[i];
    mat3_tvec(Rt, d, pd);
    log6(Rd, pd, err);
    double en = 0; for (int i = 0; i < 6; ++i) en += err[i] * err[i];
    en = sqrt(en);
    if (en < min_err) { min_err = en; memcpy(out, q, sizeof(double) * n); }
    if (en < eps) { success = 1; break; }
    if (it >= max_it) break;
    /* local-frame Jacobian: rotate both blocks by R^T, zero masked columns */
    double Jl[6 * ORC_MAXJ];
    for (int i = 0; i < n; ++i) {
      double l[3] = {J[i], J[n + i], J[2 * n + i]}, a[3] = {J[3 * n + i], J[4 * n + i], J[5 * n + i]}, lo[3], ao[3];
      mat3_tvec(R, l, lo); mat3_tvec(R, a, ao);
      const double on = (!mask || mask[i]) ? 1.0 : 0.0;
      for (int r = 0; r < 3; ++r) { Jl[r * n + i] = on * lo[r]; Jl[(3 + r) * n + i] = on * ao[r]; }
    }
    double A[36], y[6];
    for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) { double s = 0; for (int i = 0; i < n; ++i) s += Jl[r * n + i] * Jl[c * n + i]; A[6 * r + c] = s + (r == c ? damp : 0); }
    if (chol_solve6(A, err, y)) break;
    for (int i = 0; i < n; ++i) { double s = 0; for (int r = 0; r < 6; ++r) s += Jl[r * n + i] * y[r]; q[i] -= s * dt; }
  }
  if (best_err_norm) *best_err_norm = min_err;
  return success;
}

/* ---- batch drivers: one env per call like the reference's CPU path.  Single-threaded; the
 * Python binding fans chunks out over host threads (ctypes releases the GIL). ---- */
int orc_delta_step_batch(const orc_frame* fr, int nf, int n, const double* q, const double* a, int nc, const uint8_t* mask,
                         int B, int c, double* out) {
  for (int e = 0; e < B; ++e) orc_delta_step(fr, nf, q + (size_t)e * n, a + (size_t)e * nc, nc, mask, out + (size_t)e * c);
  return 0;
}
int orc_iterative_batch(const orc_frame* fr, int nf, int n, const double* q, const double* tgt, const uint8_t* mask, int B,
                        int c, int max_it, double lr, double reg, double pos_tol, double rot_tol, int patience, double* out,
                        int32_t* iters) {
  for (int e = 0; e < B; ++e) {
    int it = orc_iterative(fr, nf, q + (size_t)e * n, tgt + (size_t)e * 7, mask, max_it, lr, reg, pos_tol, rot_tol, patience,
                           out + (size_t)e * c);
    if (iters) iters[e] = it;
  }
  return 0;
}
int orc_clik_batch(const orc_frame* fr, int nf, int n, const double* q, const double* tgt, const uint8_t* mask, int B,
                   int max_it, double eps, double dt, double damp, double* out, uint8_t* success) {
  for (int e = 0; e < B; ++e) {
    int ok = orc_clik(fr, nf, q + (size_t)e * n, tgt + (size_t)e * 7, mask, max_it, eps, dt, damp, out + (size_t)e * n, 0);
    if (success) success[e] = (uint8_t)(ok == 1);
  }
  return 0;
}
