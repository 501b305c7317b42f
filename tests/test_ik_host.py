"""Path B's per-env arithmetic (csrc/msik_ik.cuh: the code target_ik_kernel runs) compiled with g++ and held to the
fp64 oracle on the CPU -- no GPU needed.

The CUDA qualifiers / intrinsics are mapped to libm by -DMSIK_HOST_TEST (csrc/msik_math.cuh), so what runs here is the
same fp32 expression tree as on the device up to MUFU roundings: the reduced mid-chain-frame system with the pose
error formed in that frame (chaingen.py _gen_reduced_ik), rot_err's polynomial / quaternion branches, the guarded
fast step and the exact damped step (dls_solve = QR of [J^T; lambda I]).  Oracle: oracle.ik_iterative
(pk.PseudoInverseIK as called at kinematics.py:144-155)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import FETCH, FLOATING, PANDA, REALMAN_R, ROOT, XARM7
from oracle import oracle as O

CHAINS = {  # struct name -> (robot fixture, c0, nr)
    "panda_tcp": (PANDA, 0, 7),
    "fetch_gripper": (FETCH, 4, 7),
    "xarm7_base": (XARM7, 0, 7),
    "realman_r_link6": (REALMAN_R, 3, 6),
    "floating_gripper_tcp": (FLOATING, 0, 6),
}

_HARNESS = r"""
#define MSIK_HOST_TEST 1
#include "../../include/msik.h"
#include "msik_math.cuh"
#include "generated/chains.cuh"
#include "msik_ik.cuh"
using namespace msik;
template <class Ch, int C0, int NR>
static int solve(const float* q0, const float* tgt, const msik_iter_cfg* it, float rtol, int masked, float* q_out, int* iters, float* err) {
  constexpr int N = Ch::N;
  msik_chain_desc d{};
  float q[N], tp[3] = {tgt[0], tgt[1], tgt[2]}, Rt[9];
  for (int i = 0; i < N; ++i) q[i] = q0[i];
  float n = std::sqrt(tgt[3] * tgt[3] + tgt[4] * tgt[4] + tgt[5] * tgt[5] + tgt[6] * tgt[6]);
  quat_to_mat(Quat{tgt[3] / n, tgt[4] / n, tgt[5] / n, tgt[6] / n}, Rt);
  const uint32_t mask = ((1u << NR) - 1u) << C0;
  IkLoop s;
  int sw;
  if (masked) sw = ik_solve_env<Ch, 0, N, true>(d, mask, *it, rtol, q, tp, Rt, s);
  else sw = ik_solve_env<Ch, C0, NR, false>(d, mask, *it, rtol, q, tp, Rt, s);
  for (int i = 0; i < N; ++i) q_out[i] = q[i];
  *iters = s.iters;
  err[0] = std::sqrt(s.epos2); err[1] = std::sqrt(s.erot2);
  return sw;
}
// rot_err alone: T, C row-major
extern "C" void rot_err_host(const float* T, const float* Cm, float* w) {
  float t[9], c[9], o[3];
  for (int i = 0; i < 9; ++i) { t[i] = T[i]; c[i] = Cm[i]; }
  rot_err(t, c, o);
  w[0] = o[0]; w[1] = o[1]; w[2] = o[2];
}
// dls_solve alone: A = J^T (7 x 6)
extern "C" void dls_host(const float* At, const float* b, float lambda, float* x) {
  float A[7][6], bb[6], xx[7];
  for (int r = 0; r < 7; ++r) for (int k = 0; k < 6; ++k) A[r][k] = At[6 * r + k];
  for (int k = 0; k < 6; ++k) bb[k] = b[k];
  dls_solve<7, 6>(A, bb, xx, lambda);
  for (int r = 0; r < 7; ++r) x[r] = xx[r];
}
extern "C" int ik_panda_tcp(const float* q0, const float* t, const msik_iter_cfg* it, float rtol, int m, float* q, int* n, float* e) {
  return solve<MsikChain_panda_tcp, 0, 7>(q0, t, it, rtol, m, q, n, e); }
extern "C" int ik_fetch_gripper(const float* q0, const float* t, const msik_iter_cfg* it, float rtol, int m, float* q, int* n, float* e) {
  return solve<MsikChain_fetch_gripper, 4, 7>(q0, t, it, rtol, m, q, n, e); }
extern "C" int ik_xarm7_base(const float* q0, const float* t, const msik_iter_cfg* it, float rtol, int m, float* q, int* n, float* e) {
  return solve<MsikChain_xarm7_base, 0, 7>(q0, t, it, rtol, m, q, n, e); }
extern "C" int ik_floating_gripper_tcp(const float* q0, const float* t, const msik_iter_cfg* it, float rtol, int m, float* q, int* n, float* e) {
  return solve<MsikChain_floating_gripper_tcp, 0, 6>(q0, t, it, rtol, m, q, n, e); }
extern "C" int ik_realman_r_link6(const float* q0, const float* t, const msik_iter_cfg* it, float rtol, int m, float* q, int* n, float* e) {
  return solve<MsikChain_realman_r_link6, 3, 6>(q0, t, it, rtol, m, q, n, e); }
"""


class IterCfg(C.Structure):  # include/msik.h msik_iter_cfg
    _fields_ = [("max_iterations", C.c_int32), ("patience", C.c_int32), ("lr", C.c_float), ("pos_tol", C.c_float),
                ("rot_tol", C.c_float), ("rank_rtol", C.c_float), ("damping", C.c_float), ("guard_pivot", C.c_float), ("mapping", C.c_int32)]


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    d = tmp_path_factory.mktemp("ikhost")
    src = d / "harness.cpp"
    src.write_text(_HARNESS)
    so = d / "harness.so"
    r = subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off", "-I" + os.path.join(ROOT, "maniskill_b200", "csrc"),
                        str(src), "-o", str(so)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return C.CDLL(str(so))


def _fp(a):
    return np.ascontiguousarray(a, dtype=np.float32).ctypes.data_as(C.POINTER(C.c_float))


def host_ik(h, name, q0, tgt, damping=1e-9, guard=0.0, masked=False, max_it=200, patience=2):
    it = IterCfg(max_it, patience, 0.2, 1e-3, 1e-2, 0.0, damping, guard, 0)
    fn = getattr(h, "ik_" + name)
    fn.restype = C.c_int
    B, n = q0.shape
    q = np.zeros((B, n), np.float32)
    iters = np.zeros(B, np.int32)
    err = np.zeros((B, 2), np.float32)
    sw = np.zeros(B, np.int32)
    q0 = np.ascontiguousarray(q0, np.float32)
    tgt = np.ascontiguousarray(tgt, np.float32)
    for e in range(B):
        k = C.c_int(0)
        sw[e] = fn(_fp(q0[e]), _fp(tgt[e]), C.byref(it), C.c_float(7 * 1.1920929e-07), int(masked), q[e].ctypes.data_as(C.POINTER(C.c_float)),
                   C.byref(k), err[e].ctypes.data_as(C.POINTER(C.c_float)))
        iters[e] = k.value
    return q, iters, err, sw


def targets_near(fr, q, rng, spread):
    q2 = q + spread * (rng.random(q.shape) - 0.5)
    _, R, p = O.jacobian(fr, q2)
    return np.concatenate([p, O.matrix_to_quaternion(R)], 1)


def test_rot_err_matches_the_quaternion_axis_angle(harness):
    """Both branches of rot_err (polynomial <= 45 deg, quaternion above) against the oracle's
    axis_angle(q_T (x) conj(q_C)) (rotation_conversions.py:528-556)."""
    from scipy.spatial.transform import Rotation as Rsc

    rng = np.random.default_rng(0)
    worst = 0.0
    for ang_max in (1e-4, 1e-2, 0.3, 0.78, 0.8, 1.5, 3.0):
        for _ in range(200):
            Cm = Rsc.random(random_state=rng.integers(1 << 30)).as_matrix()
            axis = rng.normal(size=3)
            axis /= np.linalg.norm(axis)
            ang = rng.uniform(0, ang_max)
            T = Rsc.from_rotvec(axis * ang).as_matrix() @ Cm
            w = np.zeros(3, np.float32)
            harness.rot_err_host(_fp(T.ravel()), _fp(Cm.ravel()), w.ctypes.data_as(C.POINTER(C.c_float)))
            qT = O.matrix_to_quaternion(T[None])
            qC = O.matrix_to_quaternion(Cm[None])
            ref = O.quaternion_to_axis_angle(O.quaternion_multiply(O.standardize_quaternion(qT), O.quaternion_invert(qC)))[0]
            worst = max(worst, float(np.abs(w - ref).max()))
            assert np.abs(w - ref).max() < 2e-6, (ang, w, ref)
    assert worst < 2e-6


def test_dls_solve_is_the_damped_update(harness):
    """dls_solve (QR of [J^T; lambda I]) == J^T (J J^T + lambda^2 I)^-1 b, also at an exactly singular J."""
    fr = O.load_chain(PANDA[0], PANDA[1])
    rng = np.random.default_rng(1)
    lim = np.clip(O.joint_limits(fr), -3, 3)
    qs = rng.uniform(lim[:, 0], lim[:, 1], (100, 7))
    qs[0] = 0.0  # Panda at q = 0: singular
    qs[1, 3] = -0.0698
    for lam2 in (1e-9, 1e-6, 1e-2):
        for q in qs:
            J = O.jacobian(fr, q[None])[0][0]
            b = rng.uniform(-0.2, 0.2, 6)
            ref = J.T @ np.linalg.solve(J @ J.T + lam2 * np.eye(6), b)
            x = np.zeros(7, np.float32)
            J32 = J.astype(np.float32)
            harness.dls_host(_fp(J32.T.ravel()), _fp(b), C.c_float(np.sqrt(lam2)), x.ctypes.data_as(C.POINTER(C.c_float)))
            # compare against the damped update of the SAME fp32 Jacobian (the conditioning of J itself is not under test)
            J64 = J32.astype(np.float64)
            b64 = b.astype(np.float32).astype(np.float64)
            ref32 = J64.T @ np.linalg.solve(J64 @ J64.T + lam2 * np.eye(6), b64)
            s = np.linalg.svd(J64, compute_uv=False)
            # backward-stable QR of the augmented matrix: error ~ cond(R~) eps |x|, cond(R~) = sigma_max / sqrt(sigma_min^2 + lambda^2)
            cond_aug = s.max() / np.sqrt(s.min() ** 2 + lam2)
            assert np.abs(x - ref32).max() < 8 * 6e-8 * cond_aug * np.abs(ref32).max() + 1e-6, (lam2, s.min(), np.abs(x - ref32).max())
            del ref


@pytest.mark.parametrize("name", list(CHAINS))
def test_guarded_fast_step_matches_the_damped_oracle(harness, name):
    """Default configuration (damping 1e-9, guard derived from it): iteration counts and joint targets of the fp32 loop
    against the fp64 damped-least-squares oracle; envs whose stopping iteration flips by rounding are identified as
    such (the oracle forced to the kernel's iteration count reproduces them) and must be rare."""
    robot, c0, nr = CHAINS[name]
    fr = O.load_chain(robot[0], robot[1])
    n = len(fr) if not hasattr(fr, "n") else fr.n
    lim = np.clip(O.joint_limits(fr), -3.0, 3.0)
    n = len(lim)
    rng = np.random.default_rng(3)
    B = 96
    q0 = rng.uniform(lim[:, 0] + 0.15 * (lim[:, 1] - lim[:, 0]), lim[:, 1] - 0.15 * (lim[:, 1] - lim[:, 0]), (B, n))
    if name == "fetch_gripper":
        q0[:, :2] = rng.uniform(-2, 2, (B, 2))
    mask = np.zeros(n, bool)
    mask[c0:c0 + nr] = True
    q0 = q0.astype(np.float32).astype(np.float64)
    q2 = q0.copy()
    q2[:, mask] += 0.3 * (rng.random((B, nr)) - 0.5)
    _, R, p = O.jacobian(fr, q2)
    tgt = np.concatenate([p, O.matrix_to_quaternion(R)], 1).astype(np.float32).astype(np.float64)
    Jc = O.jacobian(fr, q0)[0][:, :, mask]
    good = np.linalg.cond(Jc) < 100
    ref, it_ref = O.ik_iterative(fr, q0, tgt, col_mask=mask, return_info=True)
    got, it_got, err, sw = host_ik(harness, name, q0, tgt)
    got = got[:, mask]
    d = np.abs(got - ref).max(1)
    flip = it_got != it_ref
    # envs with the same stopping iteration agree to fp32 solve accuracy
    same = good & ~flip
    assert same.sum() > 0.8 * good.sum(), (same.sum(), good.sum())
    assert d[same].max() < 1e-5, d[same].max()
    # a flipped env is the oracle stopped at OUR iteration count
    for e in np.nonzero(good & flip)[0]:
        r2 = O.ik_iterative(fr, q0[e:e + 1], tgt[e:e + 1], col_mask=mask, max_iterations=int(it_got[e]), patience=None, pos_tol=0.0, rot_tol=0.0)
        assert np.abs(got[e] - r2[0]).max() < 1e-5, (e, it_got[e], it_ref[e])
    assert (good & flip).sum() <= max(2, 0.03 * good.sum())


def test_every_iteration_damped_equals_oracle_also_near_singularities(harness):
    """guard_pivot = 1e30: every iteration takes the exact damped step.  Near-singular starts (Panda stretched out /
    joint 4 at its limit) are where the damping acts; a larger damping makes its effect visible at fp32 level."""
    fr = O.load_chain(PANDA[0], PANDA[1])
    rng = np.random.default_rng(5)
    B = 48
    q0 = np.tile([0.0, 0.3, 0.0, -0.08, 0.0, 0.4, 0.7], (B, 1)) + 0.01 * rng.normal(size=(B, 7))  # elbow nearly straight
    q0[: B // 2] = np.asarray([0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4]) + 0.02 * rng.normal(size=(B // 2, 7))
    q0 = q0.astype(np.float32).astype(np.float64)
    tgt = targets_near(fr, q0, rng, 0.2).astype(np.float32).astype(np.float64)
    for lam2 in (1e-9, 1e-4):
        ref, it_ref = O.ik_iterative(fr, q0, tgt, reg=lam2, return_info=True)
        got, it_got, _, sw = host_ik(harness, "panda_tcp", q0, tgt, damping=lam2, guard=1e30)
        assert (sw == 0).all()  # switched to the damped step at iteration 0
        same = it_got == it_ref
        assert same.mean() > 0.9
        cond = np.linalg.cond(O.jacobian(fr, q0)[0])
        tol = 1e-5 + 2e-7 * cond  # fp32 Jacobian: cond * eps * |dq|
        assert (np.abs(got - ref).max(1)[same] < tol[same]).all(), (lam2, np.abs(got - ref).max(1)[same].max())


def test_guard_parks_near_singular_envs_and_only_those(harness):
    fr = O.load_chain(PANDA[0], PANDA[1])
    rng = np.random.default_rng(7)
    rest = np.asarray([0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4])
    B = 64
    q0 = rest + 0.02 * rng.normal(size=(B, 7))
    tgt = targets_near(fr, q0, rng, 0.3)
    _, _, _, sw = host_ik(harness, "panda_tcp", q0, tgt)
    assert (sw == -1).all(), "well-conditioned envs must stay on the fast step"
    # singular start: q = 0 (joint axes 1, 3, 5, 7 aligned) with a target a little away
    q0s = np.zeros((4, 7)) + 1e-4 * rng.normal(size=(4, 7))
    tgts = targets_near(fr, q0s, rng, 0.2)
    ref, it_ref = O.ik_iterative(fr, q0s.astype(np.float32).astype(np.float64), tgts.astype(np.float32).astype(np.float64), return_info=True)
    got, it_got, _, sw = host_ik(harness, "panda_tcp", q0s, tgts)
    assert (sw >= 0).all(), "a singular start must be handed to the damped step"
    # damping == 0: nothing parks, by definition
    _, _, _, sw0 = host_ik(harness, "panda_tcp", q0s, tgts, damping=0.0)
    assert (sw0 == -1).all()


def test_masked_dense_instantiation_matches_reduced(harness):
    """The MASKED (dense base-frame) instantiation and the reduced one solve the same problem."""
    fr = O.load_chain(PANDA[0], PANDA[1])
    rng = np.random.default_rng(9)
    rest = np.asarray([0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4])
    q0 = rest + 0.05 * rng.normal(size=(32, 7))
    tgt = targets_near(fr, q0, rng, 0.3)
    a, ia, _, _ = host_ik(harness, "panda_tcp", q0, tgt)
    b, ib, _, _ = host_ik(harness, "panda_tcp", q0, tgt, masked=True)
    same = ia == ib
    assert same.mean() > 0.9
    assert np.abs(a - b)[same].max() < 5e-6
