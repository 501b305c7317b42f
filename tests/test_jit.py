"""Run-time chain specialisation (maniskill_b200/jit.py): generated straight-line kernels for a URDF the library
has no built-in code for, compiled into a plugin and registered through the C ABI."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from maniskill_b200 import _lib as L
from maniskill_b200 import jit
from maniskill_b200.urdf import build_chain_table
from oracle import oracle as O

TOY = (os.path.join(ROOT, "tests", "golden", "toy6_kin.urdf"), "tool", [f"j{i}" for i in range(1, 7)])


def test_plugin_source_and_range_detection():
    T = build_chain_table(TOY[0], TOY[1])
    assert jit._contiguous(0x7F0) == (4, 7) and jit._contiguous(0x5) is None and jit._contiguous(0x3F) == (0, 6)
    src = jit.plugin_source("jit_x", T, 0x3C)  # joints 2..5 controlled: a contiguous range -> un-masked range kernels
    assert "struct MsikChain_jit_x" in src and "ChainRange<MsikChain_jit_x>" in src and "C0 = 2" in src and "NR = 4" in src
    assert "ChainRange<MsikChain_jit_x> {" not in jit.plugin_source("jit_x", T, 0x3F)  # full chain: no range variant
    assert jit.plugin_name(T, 0x3F) == jit.plugin_name(T) != jit.plugin_name(T, 0x3C)


def test_plugin_builds_registers_and_is_chosen_for_its_chain(built_lib):
    """nvcc cross-compiles the plugin without a GPU; registering it makes msik_chain_create pick its kernels for
    exactly this chain (host-side logic only, nothing is launched)."""
    T = build_chain_table(TOY[0], TOY[1])
    name = jit.specialise(TOY[0], TOY[1])
    assert name == jit.plugin_name(T) and os.path.exists(os.path.join(jit.PLUGIN_DIR, name + ".so"))
    lib = L.load()
    desc = L.ChainDesc()
    desc.n_joints = T.n
    packed = T.packed_f32()
    for i in range(T.n):
        desc.joint_type[i] = int(T.joint_types[i])
        desc.lower[i], desc.upper[i] = float(np.float32(T.lower[i])), float(np.float32(T.upper[i]))
    for i in range(T.n + 1):
        for k in range(12):
            desc.pre[i][k] = float(packed[i, k])
    h = C.c_void_p()
    L.check(lib.msik_chain_create(C.byref(desc), C.byref(h)))
    buf = C.create_string_buffer(64)
    L.check(lib.msik_chain_kernel_name(h, buf, 64))
    assert buf.value.decode() == name
    L.check(lib.msik_chain_destroy(h))
    desc.pre[0][9] += 0.001  # any other chain still gets the table-driven kernels
    L.check(lib.msik_chain_create(C.byref(desc), C.byref(h)))
    L.check(lib.msik_chain_kernel_name(h, buf, 64))
    assert buf.value.decode() == "generic7"
    L.check(lib.msik_chain_destroy(h))
    assert lib.msik_register_chain_ops(None) == L.MSIK_EINVAL if hasattr(L, "MSIK_EINVAL") else lib.msik_register_chain_ops(None) < 0


@pytest.mark.gpu
def test_specialised_toy_arm_matches_oracle_and_generic_kernels():
    from helpers import make_art, sample_q
    from maniskill_b200.kinematics import Kinematics
    from maniskill_b200.structs.pose import Pose

    B = 4099
    fr = O.load_chain(TOY[0], TOY[1])
    kins = {}
    for spec in (False, True):
        art = make_art(TOY, B)
        act = [j.name for j in art.active_joints]
        idx = torch.tensor([act.index(n) for n in TOY[2]], dtype=torch.int32, device="cuda")
        if not spec:  # the registry is process-wide: an earlier test may have registered this chain already
            os.environ["MSIK_FORCE_GENERIC"] = "1"
        try:
            kins[spec] = Kinematics(TOY[0], TOY[1], art, idx, specialise=spec)
        finally:
            os.environ.pop("MSIK_FORCE_GENERIC", None)
    assert kins[False].kernel_name == "generic7" and kins[True].kernel_name.startswith("jit_")
    q = sample_q(fr, B, seed=1).astype(np.float32)
    d_q = torch.from_numpy(q).cuda()
    J, R, p = O.jacobian(fr, q.astype(np.float64))
    a = np.random.default_rng(2).uniform(-0.05, 0.05, (B, 6)).astype(np.float32)
    ref = O.ik_delta_step(fr, q.astype(np.float64), a.astype(np.float64))
    cond = np.linalg.cond(J)
    ok = cond < 50
    assert ok.sum() > 500
    outs = {}
    for spec, kin in kins.items():
        pose, jac = kin.forward_kinematics(d_q, return_jacobian=True)
        assert np.abs(pose.p.cpu().numpy() - p).max() < 1e-6
        assert np.abs(jac.cpu().numpy() - J).max() < 2e-6
        out = kin.compute_ik(None, d_q, action=torch.from_numpy(a).cuda(), use_delta_ik_solver=True).cpu().numpy()
        dq = np.abs(ref - q).max(1)
        assert (np.abs(out - ref).max(1)[ok] < 1e-5 + 4e-6 * cond[ok] * np.maximum(0, dq[ok] - 0.25)).all()
        outs[spec] = out
        # iterative path: FK of the solution reaches a reachable target
        q2 = q.astype(np.float64) + np.random.default_rng(3).uniform(-0.1, 0.1, q.shape)
        _, R2, p2 = O.jacobian(fr, q2)
        tgt = np.concatenate([p2, O.matrix_to_quaternion(R2)], 1).astype(np.float32)
        sol = kin.compute_ik(Pose.create(torch.from_numpy(tgt).cuda()), d_q)
        got = kin.forward_kinematics(sol if sol.shape[1] == d_q.shape[1] else torch.cat([sol], 1))
        perr = (got.p.cpu().numpy() - tgt[:, :3])
        assert (np.linalg.norm(perr, axis=1) < 1.1e-3).mean() > 0.95
    assert np.abs(outs[True] - outs[False])[ok].max() < 1e-5
