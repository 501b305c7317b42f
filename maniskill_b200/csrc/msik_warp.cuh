// Warp-cooperative variant of path B: the mapping north_star prescribes -- one group of lanes per env, one lane per
// joint, warp shuffles for the prefix products along the chain, J J^T + lambda^2 I solved by an unrolled Cholesky.
//
//   lane j of an 8-lane group owns joint j of ONE env (4 envs per warp):
//     sincos(q_j), local transform L_j = G_j Z(q_j)                                 lane-local
//     P_j = L_0 ... L_j            inclusive scan, 3 shuffle rounds of 12 floats (Kogge-Stone, width 8)
//     EE pose = P_{n-1} G_tool     broadcast from the last joint's lane; pose error formed by every lane (SIMT: free)
//     Jacobian column c_j          lane-local from P_j and the EE position
//     S = sum_j c_j c_j^T + lambda^2 I     21 butterfly all-reduces (3 rounds), then EVERY lane factors the 6 x 6 S by
//                                  Cholesky and solves S y = dx;  dq_j = c_j . y is lane-local again
//     one step of iterative refinement with the residual taken from J itself (dx - sum_j c_j dq_j - lambda^2 y): the
//     normal equations square the condition number, the refinement brings the step back to ~cond(J) eps
//   table-driven (any chain of <= 8 joints, any controlled mask), damping exact for every value >= 0.
//
// Measured against the thread-per-env kernel (msik_kernels.cuh target_ik_kernel) on B200, profiles/r02f_ik_warp_ab.txt:
// it LOSES at every env count -- 30.7 vs 24.7 us at 256..2048 envs (pure latency: ten shuffle rounds, six dependent
// MUFU.RSQ of the Cholesky and ~770 warp instructions per iteration for 4 envs, against ~575 for 32), 39 vs 26.7 us at
// 4096, 328 vs 51 us at 65536 (it issues ~10x more warp instructions per env).  So msik_target_ik never picks it on its
// own (kIkWarpMaxEnvs = 0); it stays available as msik_iter_cfg.mapping = MSIK_IK_LANE_PER_JOINT, parity-tested, as the
// measured alternative the design was weighed against (MSIK_IK_WARP=0 / 1 and MSIK_IK_WARP_MAX_ENVS drive A/B runs).
#pragma once
#include "msik_kernels.cuh"

namespace msik {

constexpr int WLPE = 8;  // lanes per env

struct WTf {  // rigid transform: rotation (row-major) + translation
  float R[9], t[3];
};

// A * B (B expressed in A's frame)
__device__ __forceinline__ void wtf_mul(const WTf& A, const WTf& B, WTf& O) {
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      O.R[3 * r + c] = fmaf(A.R[3 * r], B.R[c], fmaf(A.R[3 * r + 1], B.R[3 + c], A.R[3 * r + 2] * B.R[6 + c]));
    O.t[r] = fmaf(A.R[3 * r], B.t[0], fmaf(A.R[3 * r + 1], B.t[1], fmaf(A.R[3 * r + 2], B.t[2], A.t[r])));
  }
}

// Cholesky factor of the symmetric positive (semi)definite S (lower triangle used), in place: L below the diagonal,
// id[k] = 1 / L[k][k].  A pivot that cancelled to rounding level (<= 5e-7 of its original diagonal entry) is dropped:
// its id = 0 zeroes that component of every solve -- the analogue of the rank cut-off of the QR solves; only reached at
// exact singularities with damping == 0.
__device__ __forceinline__ void chol6(float (&S)[6][6], float (&id)[6]) {
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const float piv = S[k][k];
    const bool ok = piv > 5e-7f * id[k];  // id[] comes in holding the original diagonal
    const float ri = ok ? rsqrt_approx(piv) : 0.f;
    id[k] = ri;
#pragma unroll
    for (int i = k + 1; i < 6; ++i) S[i][k] *= ri;
#pragma unroll
    for (int i = k + 1; i < 6; ++i)
#pragma unroll
      for (int j = k + 1; j <= i; ++j) S[i][j] = fmaf(-S[i][k], S[j][k], S[i][j]);
  }
}
__device__ __forceinline__ void chol6_solve(const float (&L)[6][6], const float (&id)[6], const float (&b)[6], float (&y)[6]) {
  float z[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float acc = b[k];
#pragma unroll
    for (int i = 0; i < k; ++i) acc = fmaf(-L[k][i], z[i], acc);
    z[k] = acc * id[k];
  }
#pragma unroll
  for (int k = 5; k >= 0; --k) {
    float acc = z[k];
#pragma unroll
    for (int i = k + 1; i < 6; ++i) acc = fmaf(-L[i][k], y[i], acc);
    y[k] = acc * id[k];
  }
}

__device__ __forceinline__ float group_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1, WLPE);
  v += __shfl_xor_sync(0xffffffffu, v, 2, WLPE);
  v += __shfl_xor_sync(0xffffffffu, v, 4, WLPE);
  return v;
}

__global__ void __launch_bounds__(TPB) target_ik_warp_kernel(const __grid_constant__ StepParams P) {
  const int tid = threadIdx.x;
  const int j = tid & (WLPE - 1);                               // this lane's joint
  const int env = (blockIdx.x * TPB + tid) / WLPE;
  const bool active = env < P.cfg.n_envs;
  const int n = P.chain.n_joints;                               // <= 8; the tool transform is pre[n]
  const bool has = j < n;
  const int ty = has ? P.chain.joint_type[j] : 0;
  const bool ctrl = has && ((P.cfg.ctrl_mask >> j) & 1u);
  const int qs = P.cfg.qpos_stride, as = P.cfg.action_stride, ts = P.cfg.target_stride;
  const int na = P.na;
  const int jj = has ? j : 0;

  float M[9], tj[3];
#pragma unroll
  for (int k = 0; k < 9; ++k) M[k] = has ? P.chain.pre[jj][k] : ((k & 3) == 0 ? 1.f : 0.f);
#pragma unroll
  for (int k = 0; k < 3; ++k) tj[k] = has ? P.chain.pre[jj][9 + k] : 0.f;
  const float* tool = P.chain.pre[n];

  const int row = (active && P.row_index) ? P.row_index[env] : env;
  float q = (active && has) ? __ldg(P.qpos + (size_t)row * qs + P.cfg.anc_col[jj]) : 0.f;
  const int rank = __popc(P.cfg.ctrl_mask & ((1u << j) - 1u));  // position among the controlled joints
  if (active && ctrl && P.sq_out) P.sq_out[(size_t)env * P.n_ctrl + rank] = q;

  // target pose (every lane of the group computes the same thing)
  float tp[3], Rt[9];
  {
    float a[6];
#pragma unroll
    for (int k = 0; k < 6; ++k)
      a[k] = (active && P.action && k < na) ? __ldg(P.action + (size_t)env * as + P.cfg.action_offset + k) : 0.f;
    float prev[7] = {0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f};
    if (active && (P.cfg.flags & (MSIK_F_USE_DELTA | MSIK_F_TARGET_GIVEN))) {
#pragma unroll
      for (int k = 0; k < 7; ++k) prev[k] = __ldg(P.prev_pose + (size_t)env * 7 + k);
    }
    Quat tq;
    if (P.cfg.flags & MSIK_F_TARGET_GIVEN) {
      tp[0] = prev[0]; tp[1] = prev[1]; tp[2] = prev[2];
      tq = Quat{prev[3], prev[4], prev[5], prev[6]};
    } else if (na == 6) {
      float b[6];
      preprocess_action<6>(P, a, b);
      compose_target<6>(P.cfg.flags, b, prev, tp, tq);
    } else {
      const float a3[3] = {a[0], a[1], a[2]};
      float b[3];
      preprocess_action<3>(P, a3, b);
      compose_target<3>(P.cfg.flags, b, prev, tp, tq);
    }
    if (active && j == 0 && P.pose_out) {
      float* o = P.pose_out + (size_t)env * 7;
      o[0] = tp[0]; o[1] = tp[1]; o[2] = tp[2];
      o[3] = tq.w; o[4] = tq.x; o[5] = tq.y; o[6] = tq.z;
    }
    const float n2 = tq.w * tq.w + tq.x * tq.x + tq.y * tq.y + tq.z * tq.z;
    const float inv = rcp_nr(sqrt_nr(n2));
    quat_to_mat(Quat{tq.w * inv, tq.x * inv, tq.y * inv, tq.z * inv}, Rt);
  }

  IkLoop s;
  ik_loop_init(s, active);
  const float damping = P.it.damping;
  for (int it = 0; it < P.it.max_iterations; ++it) {
    if (!__any_sync(0xffffffffu, s.remaining)) break;
    // local transform of this lane's joint
    WTf Pj;
    {
      float sn = 0.f, cs = 1.f;
      if (ty == MSIK_JOINT_REVOLUTE) sincos_joint(q, sn, cs);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        Pj.R[3 * r] = fmaf(cs, M[3 * r], sn * M[3 * r + 1]);
        Pj.R[3 * r + 1] = fmaf(cs, M[3 * r + 1], -sn * M[3 * r]);
        Pj.R[3 * r + 2] = M[3 * r + 2];
        Pj.t[r] = ty == MSIK_JOINT_PRISMATIC ? fmaf(M[3 * r + 2], q, tj[r]) : tj[r];
      }
    }
    // inclusive prefix product along the chain
#pragma unroll
    for (int d = 1; d < WLPE; d <<= 1) {
      WTf A, O;
#pragma unroll
      for (int k = 0; k < 9; ++k) A.R[k] = __shfl_up_sync(0xffffffffu, Pj.R[k], d, WLPE);
#pragma unroll
      for (int k = 0; k < 3; ++k) A.t[k] = __shfl_up_sync(0xffffffffu, Pj.t[k], d, WLPE);
      wtf_mul(A, Pj, O);
      if (j >= d) Pj = O;
    }
    // end-link pose from the last joint's lane
    float pe[3], Re[9];
    {
      WTf L;
#pragma unroll
      for (int k = 0; k < 9; ++k) L.R[k] = __shfl_sync(0xffffffffu, Pj.R[k], n - 1, WLPE);
#pragma unroll
      for (int k = 0; k < 3; ++k) L.t[k] = __shfl_sync(0xffffffffu, Pj.t[k], n - 1, WLPE);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        pe[r] = fmaf(L.R[3 * r], tool[9], fmaf(L.R[3 * r + 1], tool[10], fmaf(L.R[3 * r + 2], tool[11], L.t[r])));
#pragma unroll
        for (int c = 0; c < 3; ++c)
          Re[3 * r + c] = fmaf(L.R[3 * r], tool[c], fmaf(L.R[3 * r + 1], tool[3 + c], L.R[3 * r + 2] * tool[6 + c]));
      }
    }
    float dx[6];
    dx[0] = tp[0] - pe[0]; dx[1] = tp[1] - pe[1]; dx[2] = tp[2] - pe[2];
    {
      float w[3];
      rot_err(Rt, Re, w);
      dx[3] = w[0]; dx[4] = w[1]; dx[5] = w[2];
    }
    // this lane's Jacobian column (base frame, EE origin)
    float c[6];
    {
      const float zx = Pj.R[2], zy = Pj.R[5], zz = Pj.R[8];
      const float ex = pe[0] - Pj.t[0], ey = pe[1] - Pj.t[1], ez = pe[2] - Pj.t[2];
      const bool rev = ctrl && ty == MSIK_JOINT_REVOLUTE, pri = ctrl && ty == MSIK_JOINT_PRISMATIC;
      c[0] = rev ? zy * ez - zz * ey : (pri ? zx : 0.f);
      c[1] = rev ? zz * ex - zx * ez : (pri ? zy : 0.f);
      c[2] = rev ? zx * ey - zy * ex : (pri ? zz : 0.f);
      c[3] = rev ? zx : 0.f; c[4] = rev ? zy : 0.f; c[5] = rev ? zz : 0.f;
    }
    // S = J J^T + damping I on every lane
    float S[6][6], id[6];
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
      for (int b = 0; b <= a; ++b) S[a][b] = group_sum(c[a] * c[b]);
#pragma unroll
    for (int a = 0; a < 6; ++a) { S[a][a] += damping; id[a] = S[a][a]; }
    chol6(S, id);
    float y[6];
    chol6_solve(S, id, dx, y);
    float dq = fmaf(c[0], y[0], fmaf(c[1], y[1], fmaf(c[2], y[2], fmaf(c[3], y[3], fmaf(c[4], y[4], c[5] * y[5])))));
    {  // one refinement step: residual from J itself, not from the squared system
      float r[6], dy[6];
#pragma unroll
      for (int a = 0; a < 6; ++a) r[a] = dx[a] - group_sum(c[a] * dq) - damping * y[a];
      chol6_solve(S, id, r, dy);
#pragma unroll
      for (int a = 0; a < 6; ++a) y[a] += dy[a];
      dq = fmaf(c[0], y[0], fmaf(c[1], y[1], fmaf(c[2], y[2], fmaf(c[3], y[3], fmaf(c[4], y[4], c[5] * y[5])))));
    }
    if (s.remaining) {
      if (ctrl) q = fmaf(P.it.lr, dq, q);
      const float ep2 = fmaf(dx[0], dx[0], fmaf(dx[1], dx[1], dx[2] * dx[2]));
      const float er2 = fmaf(dx[3], dx[3], fmaf(dx[4], dx[4], dx[5] * dx[5]));
      ++s.iters;
      s.epos2 = ep2; s.erot2 = er2;
      const bool conv = ep2 < P.it.pos_tol * P.it.pos_tol && er2 < P.it.rot_tol * P.it.rot_tol;
      const float err = ep2 + er2;
      if (err < s.err_min) { s.err_min = err; s.noimp = 0; } else { ++s.noimp; }
      const bool stalled = P.it.patience >= 0 && s.noimp > P.it.patience;
      s.remaining = !conv && !stalled;
    }
  }
  if (!active) return;
  if (j == 0) {
    if (P.iters_out) P.iters_out[env] = s.iters;  // every step here IS the damped update: no hand-over to report
    if (P.err_out) {
      P.err_out[(size_t)env * 2] = sqrt_nr(s.epos2);
      P.err_out[(size_t)env * 2 + 1] = sqrt_nr(s.erot2);
    }
  }
  if (ctrl) {
    if (P.tq_out) P.tq_out[(size_t)env * P.n_ctrl + rank] = q;
    if (!(P.cfg.flags & MSIK_F_NO_DRIVE)) P.drive[(size_t)row * ts + P.cfg.anc_col[j]] = q;
  }
}

}  // namespace msik
