"""The oracle itself: pinned to the golden vectors generated from the reference's own
rotation_conversions.py (tools/gen_golden.py), to the known answers of SURVEY.md
Appendix A.1, and to mathematical identities.  CPU only."""
import os

import numpy as np
import pytest
import torch

from conftest import FETCH, PANDA, REALMAN_L, REALMAN_R, XARM7
from oracle import c_oracle as CO
from oracle import oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "rotation_golden.npz"))


def sample_q(frames, B, seed=0, shrink=0.1):
    lim = np.clip(O.joint_limits(frames), -3.0, 3.0)
    lo = lim[:, 0] + shrink * (lim[:, 1] - lim[:, 0])
    hi = lim[:, 1] - shrink * (lim[:, 1] - lim[:, 0])
    return np.random.default_rng(seed).uniform(lo, hi, (B, len(lim)))


def test_rotation_functions_match_reference_golden():
    assert np.allclose(O.euler_xyz_to_matrix(G["euler"]), G["euler_matrix"], atol=1e-14)
    assert np.allclose(O.matrix_to_quaternion(G["euler_matrix"]), G["euler_quat"], atol=1e-13)
    assert np.allclose(O.quaternion_raw_multiply(G["qa"], G["qb"]), G["q_raw_mul"], atol=1e-14)
    assert np.allclose(O.quaternion_multiply(G["qa"], G["qb"]), G["q_mul"], atol=1e-14)
    assert np.allclose(O.quaternion_apply(G["qa"], G["pts"]), G["q_apply"], atol=1e-13)
    assert np.allclose(O.quaternion_to_axis_angle(G["qa"]), G["q_axis_angle"], atol=1e-12)
    assert np.allclose(O.quaternion_to_matrix(G["qa"]), G["q_matrix"], atol=1e-14)
    assert np.allclose(O.matrix_to_quaternion(G["q_matrix"]), G["q_from_matrix"], atol=1e-13)


def test_pose_algebra_matches_reference_golden():
    a = np.concatenate([G["pts"], G["qa"]], 1)
    b = np.concatenate([G["pts"][::-1], G["qb"]], 1)
    m = O.pose_mul(a, b)
    assert np.allclose(m[:, 3:], G["pose_mul_q"], atol=1e-14)
    assert np.allclose(m[:, :3], G["pose_mul_p"], atol=1e-13)
    assert np.allclose(O.pose_inv(a)[:, :3], G["pose_inv_p"], atol=1e-13)
    ident = O.pose_mul(O.pose_inv(a), a)
    assert np.allclose(ident[:, :3], 0, atol=1e-12) and np.allclose(np.abs(ident[:, 3]), 1, atol=1e-12)


def test_known_answers_survey_appendix_a1():
    fr = O.load_chain(PANDA[0], PANDA[1])
    J, R, p = O.jacobian(fr, G["kat_q"])
    assert np.allclose(p, G["kat_p"], atol=2e-9)
    assert np.allclose(O.matrix_to_quaternion(R), G["kat_quat"], atol=2e-9)
    a = np.tile(G["kat_action"], (2, 1))
    dq = O.ik_delta_step(fr, G["kat_q"], a) - G["kat_q"]
    assert np.allclose(dq, G["kat_dq"], atol=2e-9)
    assert np.allclose(np.linalg.svd(J[0], compute_uv=False), G["kat_sigma_rest"], atol=1e-4)
    _, R0, p0 = O.jacobian(fr, np.zeros((1, 7)))
    assert np.allclose(p0[0], G["kat_p_zero"], atol=1e-12)
    assert np.allclose(O.matrix_to_quaternion(R0)[0], G["kat_quat_zero"], atol=1e-9)
    ff = O.load_chain(FETCH[0], FETCH[1])
    _, _, pf = O.jacobian(ff, G["kat_fetch_q"][None])
    assert np.allclose(pf[0], G["kat_fetch_p"], atol=2e-9)


@pytest.mark.parametrize("robot", [PANDA, FETCH, XARM7, REALMAN_R, REALMAN_L])
def test_jacobian_matches_finite_differences(robot):
    fr = O.load_chain(robot[0], robot[1])
    q = sample_q(fr, 8, seed=3)
    J, R, p = O.jacobian(fr, q)
    h = 1e-6
    for i in range(q.shape[1]):
        qp, qm = q.copy(), q.copy()
        qp[:, i] += h
        qm[:, i] -= h
        _, Rp, pp = O.jacobian(fr, qp)
        _, Rm, pm = O.jacobian(fr, qm)
        assert np.allclose((pp - pm) / (2 * h), J[:, :3, i], atol=1e-8)
        dR = (Rp - Rm) / (2 * h) @ R.transpose(0, 2, 1)
        w = np.stack([dR[:, 2, 1], dR[:, 0, 2], dR[:, 1, 0]], 1)
        assert np.allclose(w, J[:, 3:, i], atol=1e-8)


def test_delta_step_solves_the_linear_system():
    fr = O.load_chain(PANDA[0], PANDA[1])
    q = sample_q(fr, 64, seed=4)
    a = np.random.default_rng(5).uniform(-0.1, 0.1, (64, 6))
    qn = O.ik_delta_step(fr, q, a)
    J, _, _ = O.jacobian(fr, q)
    well = np.linalg.cond(J) < 1e4
    assert well.sum() > 32
    assert np.allclose((J @ (qn - q)[:, :, None])[well, :, 0], a[well], atol=1e-9)
    # minimum norm: dq is orthogonal to the null space of J
    _, _, Vt = np.linalg.svd(J)
    assert np.abs(np.einsum("bi,bi->b", Vt[:, -1], qn - q))[well].max() < 1e-9


def test_masked_columns_fetch():
    fr = O.load_chain(FETCH[0], FETCH[1])
    q = sample_q(fr, 16, seed=6)
    a = np.random.default_rng(7).uniform(-0.05, 0.05, (16, 6))
    mask = np.zeros(11, bool)
    mask[4:] = True
    qn = O.ik_delta_step(fr, q, a, col_mask=mask)
    assert qn.shape == (16, 7)
    J, _, _ = O.jacobian(fr, q)
    assert np.allclose((J[:, :, mask] @ (qn - q[:, mask])[:, :, None])[:, :, 0], a, atol=1e-7)


@pytest.mark.parametrize("robot,mask", [(PANDA, None), (FETCH, [0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1])])
def test_c_port_matches_numpy_oracle(robot, mask):
    fr = O.load_chain(robot[0], robot[1])
    q = sample_q(fr, 200, seed=8)
    n = q.shape[1]
    m = None if mask is None else np.array(mask, bool)
    a = np.random.default_rng(9).uniform(-0.1, 0.1, (200, 6))
    J, R, p = CO.fk_jac(fr, q[:20])
    Jo, Ro, po = O.jacobian(fr, q[:20])
    assert np.allclose(J, Jo, atol=1e-14) and np.allclose(p, po, atol=1e-14) and np.allclose(R, Ro, atol=1e-14)
    d, do = CO.delta_step(fr, q, a, mask=m), O.ik_delta_step(fr, q, a, col_mask=m)
    cond = np.linalg.cond(Jo if m is None else Jo[:, :, m])
    assert np.allclose(d, do, atol=1e-8, rtol=1e-8)
    d3, do3 = CO.delta_step(fr, q, a[:, :3], mask=m), O.ik_delta_step(fr, q, a[:, :3], pos_only=True, col_mask=m)
    assert np.allclose(d3, do3, atol=1e-8, rtol=1e-8)
    # iterative: reachable targets
    q2 = q[:64].copy()
    sel = np.ones(n, bool) if m is None else m
    q2[:, sel] += np.random.default_rng(10).uniform(-0.15, 0.15, (64, sel.sum()))
    _, R2, p2 = O.jacobian(fr, q2)
    tp = np.concatenate([p2, O.matrix_to_quaternion(R2)], 1)
    s, it = CO.iterative(fr, q[:64], tp, mask=m)
    so, ito = O.ik_iterative(fr, q[:64], tp, col_mask=m, return_info=True)
    assert (it == ito).all()
    assert np.allclose(s, so, atol=1e-8)


def test_iterative_converges_to_target():
    fr = O.load_chain(PANDA[0], PANDA[1])
    q = sample_q(fr, 64, seed=11, shrink=0.25)
    q2 = q + np.random.default_rng(12).uniform(-0.1, 0.1, q.shape)
    _, R2, p2 = O.jacobian(fr, q2)
    tp = np.concatenate([p2, O.matrix_to_quaternion(R2)], 1)
    sol, it = O.ik_iterative(fr, q, tp, return_info=True)
    _, Rs, ps = O.jacobian(fr, sol)
    assert (it < 200).all() and it.mean() > 3
    assert np.linalg.norm(ps - p2, axis=1).max() < 1e-3


def test_clik_reaches_target_and_reports_failure():
    fr = O.load_chain(PANDA[0], PANDA[1])
    q = sample_q(fr, 16, seed=13, shrink=0.25)
    q2 = q + np.random.default_rng(14).uniform(-0.05, 0.05, q.shape)
    _, R2, p2 = O.jacobian(fr, q2)
    tp = np.concatenate([p2, O.matrix_to_quaternion(R2)], 1)
    out, ok = CO.clik(fr, q, tp)
    assert ok.all()
    _, Rs, ps = O.jacobian(fr, out)
    assert np.linalg.norm(ps - p2, axis=1).max() < 1e-4
    for e in range(4):
        b, s, _ = O.ik_clik_single(fr, q[e], tp[e])
        assert s and np.allclose(b, out[e], atol=1e-10)
    far = tp.copy()
    far[:, :3] += 5.0  # unreachable: the reference then holds position (pd_ee_pose.py:110-111)
    _, ok2 = CO.clik(fr, q, far)
    assert not ok2.any()


def test_eager_torch_restatement_agrees_with_oracle():
    from baseline.eager_ref import EagerChain

    for robot in (PANDA, FETCH):
        ch = EagerChain(robot[0], robot[1], "cpu", torch.float64)
        fr = O.load_chain(robot[0], robot[1])
        q = sample_q(fr, 32, seed=15)
        J, pose = ch.jacobian(torch.from_numpy(q), ret_eef_pose=True)
        Jo, Ro, po = O.jacobian(fr, q)
        assert np.allclose(J.numpy(), Jo, atol=1e-13)
        assert np.allclose(pose[:, :3, 3].numpy(), po, atol=1e-13)
        assert np.allclose(pose[:, :3, :3].numpy(), Ro, atol=1e-13)


def test_controller_preprocessing():
    a = np.array([[2.0, -0.5, 0.25, 3.0, 0.0, 4.0], [0.1, 0.2, 0.3, 0.1, 0.2, 0.3]])
    out = O.pose_clip_and_scale(a, -0.1, 0.1, -0.1)
    assert np.allclose(out[0], [0.1, -0.05, 0.025, -0.06, 0.0, -0.08])
    assert np.allclose(out[1], [0.01, 0.02, 0.03, -0.01, -0.02, -0.03])  # rot_lower flips the sign (K9-ii)
