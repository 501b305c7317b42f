// Per-env arithmetic of path B: the iterative solver to an absolute target pose
// (pk.PseudoInverseIK.solve as called at kinematics.py:144-155).
//
//   dx = [p_tgt - p ; axis_angle(q_tgt (x) conj(q_cur))]
//   dq = J^T (J J^T + damping I)^-1 dx          q <- q + lr dq
//
// Two step functions, same contract (dq of the controlled range + squared error norms):
//   ik_iter_fast : damping -> 0 limit, dq = min-norm solution of J dq = dx.  Chains with a generated reduced system
//                  build matrix AND pose error in the frame of a mid-chain joint (Chain::reduced_ik_sys: one pair of
//                  half chains instead of FK + Jacobian + a second FK), everything else the dense Householder solve.
//                  Also returns the smallest |pivot| (>= sigma_min of the factored matrix): the caller's guard.
//   ik_iter_dls  : the damped update itself, exact for any damping > 0, by QR of the augmented [J^T; lambda I]
//                  (msik_math.cuh dls_solve), base frame, dense.
// The kernel (msik_kernels.cuh target_ik_kernel) runs the fast step while an env's guard holds and hands an env
// that comes close to a singularity -- where the damping is no longer below fp32 resolution -- to the damped step.
//
// No CUDA runtime dependence: tests/test_ik_host.py compiles this header with g++ (-DMSIK_HOST_TEST) and holds
// ik_solve_env to the fp64 oracle on the CPU.
#pragma once
#include "../../include/msik.h"
#include "msik_math.cuh"

namespace msik {

template <class Chain>
__device__ __forceinline__ void ik_sincos(const msik_chain_desc& d, const float (&q)[Chain::N], float (&sn)[Chain::N],
                                          float (&cs)[Chain::N]) {
  if constexpr (!Chain::kGeneric) {
    sincos_revolute_pairs<Chain::N, Chain::kRevoluteMask>(q, sn, cs);
    return;
  }
#pragma unroll
  for (int i = 0; i < Chain::N; ++i) {
    const bool rev = Chain::kGeneric ? d.joint_type[i] == MSIK_JOINT_REVOLUTE : ((Chain::kRevoluteMask >> i) & 1u) != 0;
    if (rev) sincos_joint(q[i], sn[i], cs[i]);
    else { sn[i] = 0.f; cs[i] = 1.f; }
  }
}

// does this instantiation take the reduced (mid-chain frame) system?
template <class Chain, int C0, int NR, bool MASKED>
struct IkUsesReduced {
#ifdef MSIK_NO_REDUCED
  static constexpr bool value = false;
#else
  static constexpr bool value = Chain::kHasReduced && !MASKED && C0 == Chain::kRedC0 && NR == Chain::kRedNR;
#endif
};

// dense base-frame system of one iteration: A = J^T rows of the controlled range (uncontrolled rows zeroed), dx
template <class Chain, int C0, int NR, bool MASKED>
__device__ __forceinline__ void ik_dense_system(const msik_chain_desc& d, uint32_t ctrl_mask, const float (&q)[Chain::N],
                                                const float (&tp)[3], const float (&Rt)[9], float (&A)[NR][6],
                                                float (&dx)[6]) {
  constexpr int N = Chain::N;
  float sn[N], cs[N];
  ik_sincos<Chain>(d, q, sn, cs);
  float Jl[N][3], Ja[N][3], pe[3], Re[9];
  Chain::fkjac(d, q, sn, cs, Jl, Ja, pe, Re);
  dx[0] = tp[0] - pe[0]; dx[1] = tp[1] - pe[1]; dx[2] = tp[2] - pe[2];
  float w[3];
  rot_err(Rt, Re, w);
  dx[3] = w[0]; dx[4] = w[1]; dx[5] = w[2];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const bool on = !MASKED || ((ctrl_mask >> (C0 + r)) & 1u);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      A[r][k] = on ? Jl[C0 + r][k] : 0.f;
      A[r][3 + k] = on ? Ja[C0 + r][k] : 0.f;
    }
  }
}

template <class Chain, int C0, int NR, bool MASKED>
__device__ __forceinline__ void ik_iter_fast(const msik_chain_desc& d, uint32_t ctrl_mask, float rtol,
                                             const float (&q)[Chain::N], const float (&tp)[3], const float (&Rt)[9],
                                             float (&dq)[NR], float& ep2, float& er2, float& pmin) {
  if constexpr (IkUsesReduced<Chain, C0, NR, MASKED>::value) {
    constexpr int N = Chain::N, NU = Chain::kRedNU;
    float sn[N], cs[N];
    ik_sincos<Chain>(d, q, sn, cs);
    float At[NU][4], Bm[2][NU], dp[3], Tm[9], Cm[9], pe[3];
    Chain::reduced_ik_sys(q, sn, cs, tp, Rt, At, Bm, dp, Tm, Cm, pe);
    float w[3];
    rot_err(Tm, Cm, w);  // norms are frame independent
    ep2 = fmaf(dp[0], dp[0], fmaf(dp[1], dp[1], dp[2] * dp[2]));
    er2 = fmaf(w[0], w[0], fmaf(w[1], w[1], w[2] * w[2]));
    float a1[4], a2[2], u[NU], v[2];
    Chain::reduced_ik_rhs(dp, w, pe, a1, a2);
    reduced_min_norm<NU, Chain::kRedSkip, Chain::kRedZB, true>(At, Bm, a1, a2, u, v, rtol, &pmin);
    Chain::reduced_scatter(u, v, dq);
  } else {
    float A[NR][6], dx[6];
    ik_dense_system<Chain, C0, NR, MASKED>(d, ctrl_mask, q, tp, Rt, A, dx);
    ep2 = fmaf(dx[0], dx[0], fmaf(dx[1], dx[1], dx[2] * dx[2]));
    er2 = fmaf(dx[3], dx[3], fmaf(dx[4], dx[4], dx[5] * dx[5]));
    min_norm_solve<NR, 6, (!MASKED && Chain::kLastLinZero && C0 + NR == Chain::N) ? 3 : 0>(A, dx, dq, rtol, &pmin);
  }
}

template <class Chain, int C0, int NR, bool MASKED>
__device__ __forceinline__ void ik_iter_dls(const msik_chain_desc& d, uint32_t ctrl_mask, float lambda,
                                            const float (&q)[Chain::N], const float (&tp)[3], const float (&Rt)[9],
                                            float (&dq)[NR], float& ep2, float& er2) {
  float A[NR][6], dx[6];
  ik_dense_system<Chain, C0, NR, MASKED>(d, ctrl_mask, q, tp, Rt, A, dx);
  ep2 = fmaf(dx[0], dx[0], fmaf(dx[1], dx[1], dx[2] * dx[2]));
  er2 = fmaf(dx[3], dx[3], fmaf(dx[4], dx[4], dx[5] * dx[5]));
  dls_solve<NR, 6>(A, dx, dq, lambda);
}

// Per-env loop state and the bookkeeping of one iteration (the same for both step functions):
// the stored solution is the iterate AFTER the update of the iteration whose pre-update error met both tolerances
// (early_stopping_any_converged), an env that has not converged stops after patience + 1 consecutive iterations
// without a new best error (per env; the upstream rule is per batch -- SURVEY A.2, documented deviation).
struct IkLoop {
  float err_min, epos2, erot2;
  int noimp, iters;
  bool remaining;
};
__device__ __forceinline__ void ik_loop_init(IkLoop& s, bool active) {
  s.err_min = 3.0e38f; s.epos2 = 0.f; s.erot2 = 0.f; s.noimp = 0; s.iters = 0; s.remaining = active;
}
template <class Chain, int C0, int NR, bool MASKED>
__device__ __forceinline__ void ik_advance(IkLoop& s, const msik_iter_cfg& it, uint32_t ctrl_mask, float (&q)[Chain::N],
                                           const float (&dq)[NR], float ep2, float er2) {
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const bool on = !MASKED || ((ctrl_mask >> (C0 + r)) & 1u);
    if (on) q[C0 + r] = fmaf(it.lr, dq[r], q[C0 + r]);
  }
  ++s.iters;
  s.epos2 = ep2; s.erot2 = er2;
  const bool conv = ep2 < it.pos_tol * it.pos_tol && er2 < it.rot_tol * it.rot_tol;
  const float err = ep2 + er2;  // squared norm of dx: same ordering as the norm
  if (err < s.err_min) { s.err_min = err; s.noimp = 0; } else { ++s.noimp; }
  const bool stalled = it.patience >= 0 && s.noimp > it.patience;
  s.remaining = !conv && !stalled;
}

// pivot guard of the fast step for a given damping (= lambda^2).  The damping changes dq by lambda^2 / sigma^2
// (relative, along each singular direction); an fp32 solve resolves it to ~ (sigma_max / sigma) eps.  The damping is
// therefore below the resolution of the fast step while sigma > lambda^2 / (c sigma_max eps) ~ 2.8e-3 for 1e-9;
// a Householder pivot can exceed the smallest singular value of a 4..6-column factor by a factor of ~8 at most.
#ifdef MSIK_HOST_TEST
inline float ik_pivot_guard(float damping) { return damping * 2.5e7f; }
#else
__host__ __device__ __forceinline__ float ik_pivot_guard(float damping) { return damping * 2.5e7f; }
#endif

// Whole solve of one env (host test entry; the kernel runs the same steps under warp-level control).
// guard < 0: fast step only (damping = 0).  Returns the iteration at which the env switched to the damped step
// (-1: never).
template <class Chain, int C0, int NR, bool MASKED>
__device__ __forceinline__ int ik_solve_env(const msik_chain_desc& d, uint32_t ctrl_mask, const msik_iter_cfg& it, float rtol,
                                            float (&q)[Chain::N], const float (&tp)[3], const float (&Rt)[9], IkLoop& s) {
  ik_loop_init(s, true);
  const float guard = it.damping > 0.f ? (it.guard_pivot > 0.f ? it.guard_pivot : ik_pivot_guard(it.damping)) : -1.f;
  const float lambda = it.damping > 0.f ? sqrtf(it.damping) : 0.f;
  bool damped = false;
  int switched = -1;
  for (int k = 0; k < it.max_iterations && s.remaining; ++k) {
    float dq[NR], ep2, er2, pmin;
    if (!damped) {
      ik_iter_fast<Chain, C0, NR, MASKED>(d, ctrl_mask, rtol, q, tp, Rt, dq, ep2, er2, pmin);
      if (pmin < guard) { damped = true; switched = k; }
    }
    if (damped) ik_iter_dls<Chain, C0, NR, MASKED>(d, ctrl_mask, lambda, q, tp, Rt, dq, ep2, er2);
    ik_advance<Chain, C0, NR, MASKED>(s, it, ctrl_mask, q, dq, ep2, er2);
  }
  return switched;
}

}  // namespace msik
