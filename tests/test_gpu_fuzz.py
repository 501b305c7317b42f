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
