"""Randomised C-ABI layouts for msik_delta_step_fused against a numpy model built from the oracle:
row strides, column placement of the chain joints, action offset, row_index gather, sibling (aux) columns,
normalisation / pos-only / no-drive flags, ragged env counts.  Whatever kernel the library picks for a layout
(persistent TMA, tiled, gather + aux pass) must write exactly the model's rows and nothing else."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import FETCH, PANDA, REALMAN_R
from helpers import PANDA_REST, make_art, sample_q
from maniskill_b200 import _lib as L
from maniskill_b200.kinematics import Kinematics
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _kin(robot, B):
    art = make_art(robot, B)
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in robot[2]], dtype=torch.int32, device="cuda")
    return Kinematics(robot[0], robot[1], art, idx)


def _case(rng, robot, kin, fr):
    n = kin.chain_table.n
    c_in_anc = np.asarray(kin.controlled_joints_idx_in_qmask)
    B = int(rng.choice([1, 31, 32, 33, 257, 1000, 2048, 4099]))
    extra = int(rng.integers(0, 6))
    qs = n + extra
    identity = rng.random() < 0.5
    anc_col = np.arange(n) if identity else np.sort(rng.choice(qs, n, replace=False))
    ts = qs if rng.random() < 0.7 else qs + int(rng.integers(1, 4))
    pos_only = rng.random() < 0.3
    nd = 3 if pos_only else 6
    aoff = int(rng.integers(0, 3))
    free_cols = [c for c in range(min(qs, ts)) if c not in set(anc_col[c_in_anc].tolist())]
    n_aux = int(rng.integers(0, min(4, len(free_cols)) + 1))
    astride = aoff + nd + int(rng.integers(0, 3)) + (1 if n_aux else 0)
    gather = rng.random() < 0.3
    no_drive = n_aux == 0 and rng.random() < 0.25
    normalize = rng.random() < 0.6
    M = B + int(rng.integers(0, 5)) if gather else B  # rows of the PhysX buffer
    rows = rng.permutation(M)[:B].astype(np.int32) if gather else None
    q_anc = sample_q(fr, B, seed=int(rng.integers(1 << 30)), rest=PANDA_REST if robot is PANDA else None, noise=0.3)
    qbuf = rng.normal(0, 1, (M, qs)).astype(np.float32)
    ridx = rows if gather else np.arange(B)
    qbuf[np.ix_(ridx, anc_col)] = q_anc.astype(np.float32)
    act = rng.uniform(-1.3, 1.3, (B, astride)).astype(np.float32)
    if not normalize:
        act[:, aoff:aoff + nd] *= 0.08
    tbuf = rng.normal(5, 1, (M, ts)).astype(np.float32)
    aux = L.AuxCfg()
    aux.n_cols = n_aux
    aux_spec = []
    uniform_aux = rng.random() < 0.5
    acols = rng.choice(free_cols, n_aux, replace=False) if n_aux else []
    for k, col in enumerate(acols):
        a_col = astride - 1 if uniform_aux else int(rng.choice([c for c in range(astride) if not aoff <= c < aoff + nd]))
        fl = (L.AUX_NORMALIZE if (rng.random() < 0.7 or uniform_aux) else 0) | (0 if uniform_aux else (L.AUX_DELTA if rng.random() < 0.5 else 0))
        mid, half = (0.02, 0.02) if uniform_aux else (float(rng.uniform(-0.5, 0.5)), float(rng.uniform(0.01, 0.5)))
        aux.col[k], aux.action_col[k], aux.flags[k], aux.mid[k], aux.half[k] = int(col), a_col, fl, mid, half
        aux_spec.append((int(col), a_col, fl, np.float32(mid), np.float32(half)))
    return dict(B=B, qs=qs, ts=ts, astride=astride, aoff=aoff, nd=nd, pos_only=pos_only, anc_col=anc_col, gather=gather, rows=rows,
                no_drive=no_drive, normalize=normalize, qbuf=qbuf, act=act, tbuf=tbuf, aux=aux, aux_spec=aux_spec, ridx=ridx)


def _model(fr, kin, cs):
    """numpy restatement of the fused step on the case's buffers (fp64 solve, fp32 sibling arithmetic)."""
    n = kin.chain_table.n
    mask = np.zeros(n, bool)
    mask[np.asarray(kin.controlled_joints_idx_in_qmask)] = True
    qa = cs["qbuf"][np.ix_(cs["ridx"], cs["anc_col"])].astype(np.float64)
    a = cs["act"][:, cs["aoff"]:cs["aoff"] + cs["nd"]].astype(np.float64)
    if cs["normalize"]:
        a = O.clip_and_scale_action(a, -0.1, 0.1) if cs["pos_only"] else O.pose_clip_and_scale(a, -0.1, 0.1, -0.1)
    new = O.ik_delta_step(fr, qa, a, pos_only=cs["pos_only"], col_mask=mask)
    out = cs["tbuf"].copy()
    ctrl_cols = cs["anc_col"][mask]
    if not cs["no_drive"]:
        out[np.ix_(cs["ridx"], ctrl_cols)] = new.astype(np.float32)
        for col, a_col, fl, mid, half in cs["aux_spec"]:
            v = cs["act"][:, a_col].copy()
            if fl & L.AUX_NORMALIZE:
                v = (mid + (half * np.clip(v, np.float32(-1), np.float32(1))).astype(np.float32)).astype(np.float32)
            if fl & L.AUX_DELTA:
                v = (cs["qbuf"][cs["ridx"], col] + v).astype(np.float32)
            out[cs["ridx"], col] = v
    J, _, _ = O.jacobian(fr, qa)
    J = J[:, :3] if cs["pos_only"] else J
    return out, new, ctrl_cols, np.linalg.cond(J[:, :, mask]), qa[:, mask]


@pytest.mark.parametrize("robot_name,seed", [("panda", s) for s in range(24)] + [("fetch", s) for s in range(100, 112)]
                         + [("realman", s) for s in range(200, 206)])
def test_random_layouts_against_the_model(robot_name, seed):
    robot = {"panda": PANDA, "fetch": FETCH, "realman": REALMAN_R}[robot_name]
    rng = np.random.default_rng(seed)
    kin = _kin(robot, 8)
    fr = O.load_chain(robot[0], robot[1])
    for _ in range(6):
        cs = _case(rng, robot, kin, fr)
        B = cs["B"]
        d_q, d_a, d_t = (torch.from_numpy(cs[k]).cuda() for k in ("qbuf", "act", "tbuf"))
        d_rows = torch.from_numpy(cs["rows"]).cuda() if cs["gather"] else None
        tq = torch.full((B, kin.n_ctrl), 7.0, device="cuda")
        sq = torch.full((B, kin.n_ctrl), 7.0, device="cuda")
        flags = L.F_USE_DELTA | (L.F_POS_ONLY if cs["pos_only"] else 0) | (L.F_NO_DRIVE if cs["no_drive"] else 0) | (L.F_NORMALIZE if cs["normalize"] else 0)
        cfg = kin.step_cfg(B, cs["qs"], cs["ts"], cs["astride"], cs["aoff"], flags, pos_low=(-0.1,) * 3, pos_high=(0.1,) * 3, rot_scale=-0.1)
        for i, c in enumerate(cs["anc_col"]):
            cfg.anc_col[i] = int(c)
        L.check(kin._lib.msik_delta_step_fused(kin._chain, C.byref(cfg), C.byref(cs["aux"]) if cs["aux"].n_cols else None, d_q.data_ptr(),
                                               d_rows.data_ptr() if d_rows is not None else None, d_a.data_ptr(),
                                               None if cs["no_drive"] else d_t.data_ptr(), tq.data_ptr(), sq.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream))
        got = d_t.cpu().numpy()
        want, new, ctrl_cols, cond, q_ctrl = _model(fr, kin, cs)
        tag = {k: cs[k] for k in ("B", "qs", "ts", "astride", "aoff", "nd", "gather", "no_drive", "normalize")} | {"aux": cs["aux_spec"], "anc": cs["anc_col"].tolist()}
        # everything but the controlled columns is bit-exact: untouched cells, padding rows, sibling columns
        other = np.ones(want.shape, bool)
        other[np.ix_(cs["ridx"], ctrl_cols)] = False
        assert np.array_equal(got[other], want[other]), tag
        ok = cond < 50
        # 1e-5 rad for controller-sized steps; a square (6-joint) arm near cond 50 can be asked for radians of
        # motion by a 0.1 m action, where the fp32 bound is relative: 4e-6 * cond * |dq| (test_gpu_kinematics)
        dq = np.abs(new - q_ctrl).max(1)
        tol = 1e-5 + 4e-6 * cond * np.maximum(0.0, dq - 0.25)
        if not cs["no_drive"]:
            err = np.abs(got[np.ix_(cs["ridx"], ctrl_cols)] - want[np.ix_(cs["ridx"], ctrl_cols)]).max(1)
            assert (err[ok] < tol[ok]).all(), (tag, err[ok].max())
        errq = np.abs(tq.cpu().numpy() - new).max(1)
        assert (errq[ok] < tol[ok]).all(), (tag, errq[ok].max())
        assert np.array_equal(sq.cpu().numpy(), q_ctrl.astype(np.float32)), tag


def _cfg(n_envs, qs, ts, astride, aoff, flags, anc_col, ctrl_mask, rot_scale=-0.1):
    cfg = L.StepCfg()
    cfg.n_envs, cfg.qpos_stride, cfg.target_stride, cfg.action_stride, cfg.action_offset = n_envs, qs, ts, astride, aoff
    cfg.flags, cfg.ctrl_mask, cfg.rot_scale = flags, ctrl_mask, rot_scale
    for i, c in enumerate(anc_col):
        cfg.anc_col[i] = int(c)
    for k in range(3):
        cfg.pos_low[k], cfg.pos_high[k] = -0.1, 0.1
    return cfg


@pytest.mark.parametrize("seed", range(8))
def test_random_layouts_state_kernels(seed):
    """msik_interp_targets / msik_reset_masked / msik_ee_pose_at_base / msik_target_pose on random strides,
    gathers and masks: pure data movement and pose algebra, so cells outside the written set are bit-exact and
    the written ones match numpy (interp / reset bit-exact, poses to 1e-6)."""
    rng = np.random.default_rng(1000 + seed)
    lib = L.load()
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(6):
        B = int(rng.choice([1, 33, 500, 4097]))
        n = int(rng.integers(6, 12))
        qs = n + int(rng.integers(0, 5))
        ts = qs + int(rng.integers(0, 3))
        anc_col = np.sort(rng.choice(qs, n, replace=False))
        ctrl = np.sort(rng.choice(n, int(rng.integers(3, n + 1)), replace=False))
        mask = int(sum(1 << int(i) for i in ctrl))
        c = len(ctrl)
        cols = anc_col[ctrl]
        gather = rng.random() < 0.5
        M = B + int(rng.integers(0, 4))
        ridx = rng.permutation(M)[:B].astype(np.int32) if gather else np.arange(B, dtype=np.int32)
        d_rows = torch.from_numpy(ridx).cuda() if gather else None
        rptr = d_rows.data_ptr() if gather else None
        cfg = _cfg(B, qs, ts, 0, 0, 0, anc_col, mask)

        # (1) interpolation step: drive[row, col_j] = fl(start + step_size * k)
        start = rng.normal(0, 1, (B, c)).astype(np.float32)
        ssz = rng.normal(0, 0.01, (B, c)).astype(np.float32)
        tbuf = rng.normal(5, 1, (M, ts)).astype(np.float32)
        d_t = torch.from_numpy(tbuf).cuda()
        k = float(rng.integers(1, 6))
        d_start, d_ssz = torch.from_numpy(start).cuda(), torch.from_numpy(ssz).cuda()  # keep the device copies alive over the call
        L.check(lib.msik_interp_targets(C.byref(cfg), d_start.data_ptr(), d_ssz.data_ptr(), C.c_float(k), rptr, d_t.data_ptr(), st))
        want = tbuf.copy()
        want[np.ix_(ridx, cols)] = (start.astype(np.float64) + ssz.astype(np.float64) * k).astype(np.float32)  # one fma rounding
        got = d_t.cpu().numpy()
        other = np.ones(want.shape, bool)
        other[np.ix_(ridx, cols)] = False
        assert np.array_equal(got[other], want[other])
        assert np.abs(got - want).max() <= 2.4e-7 * 8  # fma vs mul+add: an ulp of a value of magnitude < 8

        # (2) masked reset
        qbuf = rng.normal(0, 1, (M, qs)).astype(np.float32)
        rmask = (rng.random(B) < 0.4).astype(np.uint8)
        ee = rng.normal(0, 1, (B, 7)).astype(np.float32)
        sq0, tq0, tp0 = (rng.normal(3, 1, s).astype(np.float32) for s in ((B, c), (B, c), (B, 7)))
        d_sq, d_tq, d_tp = (torch.from_numpy(x.copy()).cuda() for x in (sq0, tq0, tp0))
        d_m, d_q, d_ee = torch.from_numpy(rmask).cuda(), torch.from_numpy(qbuf).cuda(), torch.from_numpy(ee).cuda()
        L.check(lib.msik_reset_masked(C.byref(cfg), d_m.data_ptr(), d_q.data_ptr(), rptr, d_ee.data_ptr(), d_sq.data_ptr(), d_tq.data_ptr(),
                                      d_tp.data_ptr(), st))
        m = rmask.astype(bool)
        cur = qbuf[np.ix_(ridx, cols)]
        assert np.array_equal(d_sq.cpu().numpy(), np.where(m[:, None], cur, sq0))
        assert np.array_equal(d_tq.cpu().numpy(), np.where(m[:, None], cur, tq0))
        assert np.array_equal(d_tp.cpu().numpy(), np.where(m[:, None], ee, tp0))

        # (3) ee_pose_at_base from rigid-body rows
        nb = 3 * B + 5
        bs = 7 + int(rng.integers(0, 7))
        body = rng.normal(0, 1, (nb, bs)).astype(np.float32)
        body[:, 3:7] /= np.linalg.norm(body[:, 3:7], axis=1, keepdims=True)
        ri, ei = rng.integers(0, nb, B).astype(np.int32), rng.integers(0, nb, B).astype(np.int32)
        d_pose = torch.full((B, 7), 9.0, device="cuda")
        d_body, d_ri, d_ei = torch.from_numpy(body).cuda(), torch.from_numpy(ri).cuda(), torch.from_numpy(ei).cuda()
        L.check(lib.msik_ee_pose_at_base(d_body.data_ptr(), bs, d_ri.data_ptr(), d_ei.data_ptr(), B, d_pose.data_ptr(), st))
        ref = O.ee_pose_at_base(body[ri, :7].astype(np.float64), body[ei, :7].astype(np.float64))
        gp = d_pose.cpu().numpy()
        assert np.abs(gp[:, :3] - ref[:, :3]).max() < 2e-6
        assert np.minimum(np.abs(gp[:, 3:] - ref[:, 3:]).max(1), np.abs(gp[:, 3:] + ref[:, 3:]).max(1)).max() < 1e-6

        # (4) compute_target_pose incl. action pre-processing, every frame combination
        pos_only = rng.random() < 0.3
        nd = 3 if pos_only else 6
        aoff = int(rng.integers(0, 3))
        astride = aoff + nd + int(rng.integers(0, 3))
        use_delta = rng.random() < 0.8
        body_t = use_delta and rng.random() < 0.5
        body_r = use_delta and (not pos_only) and rng.random() < 0.5
        flags = L.F_NORMALIZE | (L.F_POS_ONLY if pos_only else 0) | (L.F_USE_DELTA if use_delta else 0) | (L.F_BODY_TRANS if body_t else 0) | (L.F_BODY_ROT if body_r else 0)
        act = rng.uniform(-1.2, 1.2, (B, astride)).astype(np.float32)
        prev = rng.normal(0, 1, (B, 7)).astype(np.float32)
        prev[:, 3:] /= np.linalg.norm(prev[:, 3:], axis=1, keepdims=True)
        prev[:, 3:] = O.standardize_quaternion(prev[:, 3:].astype(np.float64)).astype(np.float32)
        cfg2 = _cfg(B, qs, ts, astride, aoff, flags, anc_col, mask)
        d_out = torch.full((B, 7), 9.0, device="cuda")
        d_act, d_prev = torch.from_numpy(act).cuda(), torch.from_numpy(prev).cuda()
        L.check(lib.msik_target_pose(C.byref(cfg2), d_act.data_ptr(), d_prev.data_ptr(), d_out.data_ptr(), st))
        a = act[:, aoff:aoff + nd].astype(np.float64)
        a = O.clip_and_scale_action(a, -0.1, 0.1) if pos_only else O.pose_clip_and_scale(a, -0.1, 0.1, -0.1)
        if pos_only:
            frame = "body_translation" if body_t else "root_translation"
        else:
            frame = ("body_translation" if body_t else "root_translation") + ":" + ("body_aligned_body_rotation" if body_r else "root_aligned_body_rotation")
        ref = O.compute_target_pose(prev.astype(np.float64), a, use_delta, frame, pos_only=pos_only)
        go = d_out.cpu().numpy()
        assert np.abs(go[:, :3] - ref[:, :3]).max() < 1e-6, (frame, use_delta)
        assert np.minimum(np.abs(go[:, 3:] - ref[:, 3:]).max(1), np.abs(go[:, 3:] + ref[:, 3:]).max(1)).max() < 1e-6, (frame, use_delta)


@pytest.mark.parametrize("robot_name,seed", [("panda", s) for s in range(300, 306)] + [("fetch", s) for s in range(400, 403)])
def test_random_layouts_iterative_path(robot_name, seed):
    """msik_target_ik (path B) on random layouts: absolute targets (pose given as xyz + euler XYZ action), delta
    targets from a previous pose, and given target poses.  Bystander cells are bit-exact; target_pose_out matches
    the oracle's compute_target_pose; the returned joints reach that pose within the solver's stopping tolerances
    wherever the oracle's own run of the same algorithm converged."""
    robot = PANDA if robot_name == "panda" else FETCH
    rng = np.random.default_rng(seed)
    kin = _kin(robot, 8)
    fr = O.load_chain(robot[0], robot[1])
    n = kin.chain_table.n
    c_in_anc = np.asarray(kin.controlled_joints_idx_in_qmask)
    mask = np.zeros(n, bool)
    mask[c_in_anc] = True
    for _ in range(4):
        B = int(rng.choice([1, 33, 300, 1025]))
        qs = n + int(rng.integers(0, 5))
        ts = qs + int(rng.integers(0, 2))
        anc_col = np.arange(n) if rng.random() < 0.5 else np.sort(rng.choice(qs, n, replace=False))
        gather = rng.random() < 0.4
        M = B + int(rng.integers(0, 4))
        ridx = rng.permutation(M)[:B].astype(np.int32) if gather else np.arange(B, dtype=np.int32)
        d_rows = torch.from_numpy(ridx).cuda() if gather else None
        mode = str(rng.choice(["absolute", "delta", "given"]))
        q_anc = sample_q(fr, B, seed=int(rng.integers(1 << 30)), rest=PANDA_REST if robot is PANDA else None, noise=0.2).astype(np.float32).astype(np.float64)
        qbuf = rng.normal(0, 1, (M, qs)).astype(np.float32)
        qbuf[np.ix_(ridx, anc_col)] = q_anc.astype(np.float32)
        tbuf = rng.normal(5, 1, (M, ts)).astype(np.float32)
        # a reachable goal: FK of a perturbed configuration
        q_goal = q_anc.copy()
        q_goal[:, mask] += rng.uniform(-0.12, 0.12, (B, int(mask.sum())))
        _, Rg, pg = O.jacobian(fr, q_goal)
        goal = np.concatenate([pg, O.matrix_to_quaternion(Rg)], 1)
        _, R0, p0 = O.jacobian(fr, q_anc)
        cur = np.concatenate([p0, O.matrix_to_quaternion(R0)], 1).astype(np.float32)
        aoff = int(rng.integers(0, 3))
        astride = aoff + 6 + int(rng.integers(0, 2))
        act = rng.uniform(-1, 1, (B, astride)).astype(np.float32)
        flags, prev, want_tp = 0, None, None
        if mode == "absolute":  # pd_ee_pose: action = (xyz, euler XYZ) of the goal, not normalised
            from scipy.spatial.transform import Rotation

            eul = Rotation.from_matrix(Rg.reshape(-1, 3, 3)).as_euler("XYZ")  # intrinsic: R = Rx Ry Rz (rotation_conversions.py:197-223)
            act[:, aoff:aoff + 3], act[:, aoff + 3:aoff + 6] = pg, eul
            want_tp = O.compute_target_pose(None, act[:, aoff:aoff + 6].astype(np.float64), False, "root_translation:root_aligned_body_rotation")
        if mode == "delta":  # pd_ee_target_delta_pose: previous target pose + normalised delta action
            flags = L.F_USE_DELTA | L.F_USE_TARGET | L.F_NORMALIZE
            prev = cur.copy()
            a = O.pose_clip_and_scale(act[:, aoff:aoff + 6].astype(np.float64), -0.1, 0.1, -0.1)
            want_tp = O.compute_target_pose(prev.astype(np.float64), a, True, "root_translation:root_aligned_body_rotation")
        if mode == "given":
            flags = L.F_TARGET_GIVEN
            prev = goal.astype(np.float32)
            want_tp = prev.astype(np.float64)
        cfg = kin.step_cfg(B, qs, ts, astride, aoff, flags, pos_low=(-0.1,) * 3, pos_high=(0.1,) * 3, rot_scale=-0.1)
        for i, c in enumerate(anc_col):
            cfg.anc_col[i] = int(c)
        it = kin.iter_cfg()
        d_q, d_a, d_t = torch.from_numpy(qbuf).cuda(), torch.from_numpy(act).cuda(), torch.from_numpy(tbuf).cuda()
        d_prev = torch.from_numpy(prev).cuda() if prev is not None else None
        d_tp = torch.full((B, 7), 9.0, device="cuda")
        d_tq = torch.full((B, kin.n_ctrl), 9.0, device="cuda")
        d_sq = torch.full((B, kin.n_ctrl), 9.0, device="cuda")
        d_it = torch.full((B,), -1, dtype=torch.int32, device="cuda")
        d_err = torch.full((B, 2), -1.0, device="cuda")
        L.check(kin._lib.msik_target_ik(kin._chain, C.byref(cfg), C.byref(it), d_q.data_ptr(), d_rows.data_ptr() if gather else None, d_a.data_ptr(),
                                        d_prev.data_ptr() if d_prev is not None else None, d_tp.data_ptr(), d_t.data_ptr(), d_tq.data_ptr(),
                                        d_sq.data_ptr(), d_it.data_ptr(), d_err.data_ptr(), torch.cuda.current_stream().cuda_stream))
        tag = dict(mode=mode, B=B, qs=qs, ts=ts, gather=gather, anc=anc_col.tolist(), aoff=aoff, astride=astride)
        got = d_t.cpu().numpy()
        ctrl_cols = anc_col[mask]
        other = np.ones(got.shape, bool)
        other[np.ix_(ridx, ctrl_cols)] = False
        assert np.array_equal(got[other], tbuf[other]), tag
        assert np.array_equal(d_sq.cpu().numpy(), q_anc[:, mask].astype(np.float32)), tag
        tp = d_tp.cpu().numpy()
        assert np.abs(tp[:, :3] - want_tp[:, :3]).max() < 1e-6, tag
        assert np.minimum(np.abs(tp[:, 3:] - want_tp[:, 3:]).max(1), np.abs(tp[:, 3:] + want_tp[:, 3:]).max(1)).max() < 1e-6, tag
        sol = d_tq.cpu().numpy()
        assert np.array_equal(got[np.ix_(ridx, ctrl_cols)], sol), tag
        # the oracle's run of the same algorithm, and where it converged the kernel must have reached the pose too
        ref, info = O.ik_iterative(fr, q_anc, want_tp, col_mask=mask, return_info=True)
        q_full = q_anc.copy()
        q_full[:, mask] = sol
        _, Rs, ps = O.jacobian(fr, q_full)
        perr = np.linalg.norm(ps - want_tp[:, :3], axis=1)
        qs_ = O.matrix_to_quaternion(Rs)
        rerr = 2 * np.arccos(np.clip(np.abs((qs_ * O.standardize_quaternion(want_tp[:, 3:])).sum(1)), 0, 1))
        q_ref = q_anc.copy()
        q_ref[:, mask] = ref
        _, Rr, pr = O.jacobian(fr, q_ref)
        conv = (np.linalg.norm(pr - want_tp[:, :3], axis=1) < 1e-3) & (2 * np.arccos(np.clip(np.abs((O.matrix_to_quaternion(Rr) * O.standardize_quaternion(want_tp[:, 3:])).sum(1)), 0, 1)) < 1e-2)
        assert conv.mean() > 0.5, (tag, conv.mean())  # Fetch: the algorithm itself stalls near singular samples (oracle too)
        assert (perr[conv] < 1.1e-3).all() and (rerr[conv] < 1.1e-2).all(), (tag, perr[conv].max(), rerr[conv].max())


@pytest.mark.parametrize("robot_name,seed", [("panda", 500), ("panda", 501), ("fetch", 502), ("realman", 503)])
def test_random_layouts_fk_and_host_pipeline(robot_name, seed):
    """msik_fk / msik_fk_world with gathers and scattered joint columns; msik_delta_step_host with random strides,
    offsets, chunk sizes and flags against the resident msik_delta_step on the same rows (bit-exact)."""
    robot = {"panda": PANDA, "fetch": FETCH, "realman": REALMAN_R}[robot_name]
    rng = np.random.default_rng(seed)
    kin = _kin(robot, 8)
    fr = O.load_chain(robot[0], robot[1])
    n = kin.chain_table.n
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(4):
        B = int(rng.choice([1, 33, 700, 5000]))
        qs = n + int(rng.integers(0, 5))
        anc_col = np.arange(n) if rng.random() < 0.4 else np.sort(rng.choice(qs, n, replace=False))
        gather = rng.random() < 0.5
        M = B + int(rng.integers(0, 4))
        ridx = rng.permutation(M)[:B].astype(np.int32) if gather else np.arange(B, dtype=np.int32)
        d_rows = torch.from_numpy(ridx).cuda() if gather else None
        q_anc = sample_q(fr, B, seed=int(rng.integers(1 << 30))).astype(np.float32)
        if robot_name != "panda":
            q_anc[:, :2] = rng.uniform(-1, 1, (B, 2))  # planar base near the origin (1e-6 m bar)
        qbuf = rng.normal(0, 1, (M, qs)).astype(np.float32)
        qbuf[np.ix_(ridx, anc_col)] = q_anc
        d_q = torch.from_numpy(qbuf).cuda()
        cfg = kin.step_cfg(B, qs, qs, 0, 0, 0)
        for i, c in enumerate(anc_col):
            cfg.anc_col[i] = int(c)
        pose = torch.full((B, 7), 9.0, device="cuda")
        jac = torch.full((B, 6, n), 9.0, device="cuda")
        L.check(kin._lib.msik_fk(kin._chain, C.byref(cfg), d_q.data_ptr(), d_rows.data_ptr() if gather else None, pose.data_ptr(),
                                 jac.data_ptr() if rng.random() < 0.7 else None, st))
        J, R, p = O.jacobian(fr, q_anc.astype(np.float64))
        gp = pose.cpu().numpy()
        assert np.abs(gp[:, :3] - p).max() < 2e-6
        qo = O.matrix_to_quaternion(R)
        assert np.minimum(np.abs(gp[:, 3:] - qo).max(1), np.abs(gp[:, 3:] + qo).max(1)).max() < 1e-6
        gj = jac.cpu().numpy()
        assert (gj == 9.0).all() or np.abs(gj - J).max() < 3e-6
        root = np.concatenate([rng.uniform(-1, 1, (B, 3)), O.standardize_quaternion(rng.normal(size=(B, 4)) / 1.0)], 1)
        root[:, 3:] /= np.linalg.norm(root[:, 3:], axis=1, keepdims=True)
        root = root.astype(np.float32)
        d_root = torch.from_numpy(root).cuda()
        wpose = torch.full((B, 7), 9.0, device="cuda")
        L.check(kin._lib.msik_fk_world(kin._chain, C.byref(cfg), d_q.data_ptr(), d_rows.data_ptr() if gather else None, d_root.data_ptr(),
                                       wpose.data_ptr(), st))
        ref = O.pose_mul(root.astype(np.float64), np.concatenate([p, qo], 1))
        gw = wpose.cpu().numpy()
        assert np.abs(gw[:, :3] - ref[:, :3]).max() < 3e-6
        assert np.minimum(np.abs(gw[:, 3:] - ref[:, 3:]).max(1), np.abs(gw[:, 3:] + ref[:, 3:]).max(1)).max() < 1e-6

        # host pipeline == resident path on the same rows
        pos_only = rng.random() < 0.3
        nd = 3 if pos_only else 6
        aoff = int(rng.integers(0, 3))
        astride = aoff + nd + int(rng.integers(0, 3))
        normalize = rng.random() < 0.6
        h_q = torch.from_numpy(qbuf[:B].copy()).pin_memory()  # dense rows 0..B-1 (the host entry takes no row_index)
        h_q[:, torch.from_numpy(anc_col).long()] = torch.from_numpy(q_anc)
        h_a = torch.from_numpy(rng.uniform(-1.2, 1.2, (B, astride)).astype(np.float32) * (1.0 if normalize else 0.08)).pin_memory()
        flags = L.F_USE_DELTA | L.F_NO_DRIVE | (L.F_POS_ONLY if pos_only else 0) | (L.F_NORMALIZE if normalize else 0)
        hcfg = kin.step_cfg(B, qs, qs, astride, aoff, flags, pos_low=(-0.1,) * 3, pos_high=(0.1,) * 3, rot_scale=-0.1)
        for i, c in enumerate(anc_col):
            hcfg.anc_col[i] = int(c)
        chunk = int(rng.choice([64, 1000, 4096]))
        pipe = C.c_void_p()
        L.check(kin._lib.msik_pipeline_create(chunk, qs, astride, C.byref(pipe)))
        h_o = torch.full((B, kin.n_ctrl), 9.0).pin_memory()
        try:
            L.check(kin._lib.msik_delta_step_host(pipe, kin._chain, C.byref(hcfg), h_q.data_ptr(), h_a.data_ptr(), h_o.data_ptr()))
        finally:
            L.check(kin._lib.msik_pipeline_destroy(pipe))
        dq2, da2 = h_q.cuda(), h_a.cuda()
        out = torch.full((B, kin.n_ctrl), 9.0, device="cuda")
        L.check(kin._lib.msik_delta_step(kin._chain, C.byref(hcfg), dq2.data_ptr(), None, da2.data_ptr(), None, out.data_ptr(), None, st))
        assert torch.equal(h_o, out.cpu()), dict(B=B, qs=qs, astride=astride, aoff=aoff, chunk=chunk, pos_only=pos_only)


@pytest.mark.parametrize("seed", list(range(300, 312)))
def test_dual_step_random_layouts_against_two_separate_steps(seed):
    """msik_delta_step_dual (both Realman arms in one launch where the layout allows it, the second arm's own launch
    otherwise) on random layouts -- column shift inside wider rows, target stride, action offsets of both arms, sibling
    columns, pos-only / pose, ragged env counts, row_index gathers -- against msik_delta_step_fused (arm 1 + siblings)
    followed by msik_delta_step (arm 2), the path every other test pins.  Cells no controller owns stay bit-identical."""
    from conftest import REALMAN_L

    rng = np.random.default_rng(seed)
    kr, kl = _kin(REALMAN_R, 8), _kin(REALMAN_L, 8)
    st = torch.cuda.current_stream().cuda_stream
    lib = kr._lib
    for _ in range(5):
        B = int(rng.choice([1, 31, 33, 257, 1000, 4099]))
        shift = int(rng.integers(0, 4))
        qs = 41 + shift + int(rng.integers(0, 3))
        ts = qs if rng.random() < 0.7 else qs + int(rng.integers(1, 4))
        pos_only = rng.random() < 0.5
        nd = 3 if pos_only else 6
        normalize = rng.random() < 0.7
        gather = rng.random() < 0.25
        off1 = int(rng.integers(0, 3))
        off2 = off1 + nd + int(rng.integers(0, 3))
        n_aux = int(rng.integers(0, 5))
        astride = off2 + nd + int(rng.integers(0, 2)) + n_aux
        M = B + int(rng.integers(0, 5)) if gather else B
        rows = torch.from_numpy(rng.permutation(M)[:B].astype(np.int32)).cuda() if gather else None
        flags = L.F_USE_DELTA | (L.F_POS_ONLY if pos_only else 0) | (L.F_NORMALIZE if normalize else 0)
        kw = dict(pos_low=(-0.1,) * 3, pos_high=(0.1,) * 3, rot_scale=-0.1, col_shift=shift)
        c1 = kr.step_cfg(B, qs, ts, astride, off1, flags, **kw)
        c2 = kl.step_cfg(B, qs, ts, astride, off2, flags, **kw)
        owned = {c1.anc_col[i] for i in range(3, 9)} | {c2.anc_col[i] for i in range(3, 9)}
        free = [c for c in range(min(qs, ts)) if c not in owned]
        aux = L.AuxCfg()
        aux.n_cols = n_aux
        for k, col in enumerate(rng.choice(free, n_aux, replace=False) if n_aux else []):
            aux.col[k], aux.action_col[k] = int(col), off2 + nd + int(rng.integers(0, max(astride - off2 - nd, 1)))
            aux.flags[k] = (L.AUX_NORMALIZE if rng.random() < 0.7 else 0) | (L.AUX_DELTA if rng.random() < 0.5 else 0)
            aux.mid[k], aux.half[k] = float(rng.uniform(-0.5, 0.5)), float(rng.uniform(0.01, 0.5))
        q = torch.from_numpy(rng.uniform(-1.5, 1.5, (M, qs)).astype(np.float32)).cuda()
        a = torch.from_numpy((rng.uniform(-1.3, 1.3, (B, astride)) * (1.0 if normalize else 0.08)).astype(np.float32)).cuda()
        t0 = torch.from_numpy(rng.normal(5, 1, (M, ts)).astype(np.float32)).cuda()
        auxp = C.byref(aux) if n_aux else None
        rp = rows.data_ptr() if rows is not None else None
        want = t0.clone()
        L.check(lib.msik_delta_step_fused(kr._chain, C.byref(c1), auxp, q.data_ptr(), rp, a.data_ptr(), want.data_ptr(), None, None, st))
        L.check(lib.msik_delta_step(kl._chain, C.byref(c2), q.data_ptr(), rp, a.data_ptr(), want.data_ptr(), None, None, st))
        got = t0.clone()
        d = L.DualCfg()
        d.chain, d.cfg = kl._chain, c2
        n0 = lib.msik_launch_count()
        L.check(lib.msik_delta_step_dual(kr._chain, C.byref(c1), C.byref(d), auxp, None, q.data_ptr(), rp, a.data_ptr(), got.data_ptr(), None, st))
        torch.cuda.synchronize()
        nl = lib.msik_launch_count() - n0
        tag = dict(B=B, shift=shift, qs=qs, ts=ts, pos_only=pos_only, normalize=normalize, gather=gather, off=(off1, off2), astride=astride,
                   n_aux=n_aux, launches=nl)
        aligned = (qs * 4 * 32) % 16 == 0  # always true: a 32-row span is a multiple of 16 bytes
        if not gather and aligned:
            assert nl == 1, tag
        cols = sorted(owned | {aux.col[k] for k in range(n_aux)})
        other = [c for c in range(ts) if c not in cols]
        assert torch.equal(got[:, other], t0[:, other]), tag
        assert (got - want).abs().max().item() <= 2e-6, tag
        acols = [aux.col[k] for k in range(n_aux)]
        assert torch.equal(got[:, acols], want[:, acols]), tag


def test_dual_step_rejects_inconsistent_second_controller():
    from conftest import REALMAN_L

    kr, kl = _kin(REALMAN_R, 8), _kin(REALMAN_L, 8)
    lib = kr._lib
    B, qs = 64, 41
    flags = L.F_USE_DELTA
    buf = torch.zeros(B, 64, device="cuda")
    c1 = kr.step_cfg(B, qs, qs, 12, 0, flags)
    st = torch.cuda.current_stream().cuda_stream

    def call(c2, chain2, aux=None):
        d = L.DualCfg()
        d.chain, d.cfg = chain2, c2
        return lib.msik_delta_step_dual(kr._chain, C.byref(c1), C.byref(d), C.byref(aux) if aux is not None else None, None, buf.data_ptr(), None,
                                        buf.data_ptr(), buf.data_ptr(), None, st)

    assert call(kl.step_cfg(B, qs, qs, 12, 6, flags), kl._chain) == 0
    assert call(kl.step_cfg(B, qs + 1, qs, 12, 6, flags), kl._chain) == -22  # other row stride
    assert call(kl.step_cfg(B, qs, qs, 12, 6, flags | L.F_USE_TARGET), kl._chain) == -22  # not the one-step delta path
    assert call(kl.step_cfg(B, qs, qs, 12, 8, flags), kl._chain) == -22  # action slice outside the row
    assert call(kr.step_cfg(B, qs, qs, 12, 6, flags), kr._chain) == -22  # the same arm twice: common columns
    assert "common column" in lib.msik_last_error().decode()
    aux = L.AuxCfg()
    aux.n_cols, aux.col[0], aux.action_col[0] = 1, kl.step_cfg(B, qs, qs, 12, 6, flags).anc_col[5], 0
    assert call(kl.step_cfg(B, qs, qs, 12, 6, flags), kl._chain, aux) == -22  # a sibling column the second arm drives
