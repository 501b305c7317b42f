"""Host logic that needs no GPU: URDF flattening, Pose, spaces, sharding maths, and the
C ABI (library loads, exports every symbol include/msik.h declares, argument checks)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from conftest import ASSETS, FETCH, PANDA, ROOT, XARM7
from maniskill_b200 import _lib as L
from maniskill_b200 import urdf as U
from maniskill_b200.sharding import ShardStats, reduce_stats, shard_range
from maniskill_b200.structs.pose import Pose
from oracle import oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "rotation_golden.npz"))


def fk_table(T: U.ChainTable, q):
    R, p = np.eye(3), np.zeros(3)
    for i in range(T.n):
        p = p + R @ T.t[i]
        R = R @ T.M[i]
        if T.joint_types[i] == U.JOINT_REVOLUTE:
            c, s = np.cos(q[i]), np.sin(q[i])
            R = R @ np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
        else:
            p = p + R[:, 2] * q[i]
    return R @ T.M_tool, p + R @ T.t_tool


@pytest.mark.parametrize("robot", [PANDA, FETCH, XARM7])
def test_flattened_chain_reproduces_oracle_fk(robot):
    T = U.build_chain_table(robot[0], robot[1])
    fr = O.load_chain(robot[0], robot[1])
    rng = np.random.default_rng(0)
    for _ in range(10):
        q = rng.uniform(-2, 2, T.n)
        R, p = fk_table(T, q)
        _, Ro, po = O.jacobian(fr, q[None])
        assert np.allclose(R, Ro[0], atol=1e-14) and np.allclose(p, po[0], atol=1e-14)


def test_panda_table_is_signed_permutations():
    T = U.build_chain_table(PANDA[0], PANDA[1])
    assert T.n == 7 and T.joint_names == [f"panda_joint{i}" for i in range(1, 8)]
    assert set(np.unique(np.abs(T.M))) <= {0.0, 1.0}
    assert np.allclose(T.t[0], [0, 0, 0.333]) and np.allclose(T.t_tool, [0, 0, 0.2104])
    assert np.allclose(T.lower[3], -3.0718) and np.allclose(T.upper[5], 3.7525)
    F = U.build_chain_table(FETCH[0], FETCH[1])
    assert F.n == 11 and list(F.joint_types[:4]) == [2, 2, 1, 2]
    assert set(np.unique(np.abs(F.M))) <= {0.0, 1.0}
    assert F.packed_f32().shape == (12, 12)


def test_urdf_errors():
    with pytest.raises(KeyError):
        U.build_chain_table(PANDA[0], "no_such_link")
    with pytest.raises(ValueError):
        U.build_chain_table(PANDA[0], "panda_link0")  # root: no movable joints


def test_pose_matches_reference_golden():
    a = Pose.create(torch.from_numpy(np.concatenate([G["pts"], G["qa"]], 1)))
    b = Pose.create(torch.from_numpy(np.concatenate([G["pts"][::-1].copy(), G["qb"]], 1)))
    m = a * b
    assert np.allclose(m.q.numpy(), G["pose_mul_q"], atol=1e-6)
    assert np.allclose(m.p.numpy(), G["pose_mul_p"], atol=1e-5)
    assert np.allclose(a.inv().p.numpy(), G["pose_inv_p"], atol=1e-5)
    assert a.raw_pose.dtype == torch.float32  # to_tensor semantics: f64 -> f32
    assert Pose.create_from_pq(p=[1, 2, 3]).raw_pose.tolist() == [[1, 2, 3, 1, 0, 0, 0]]
    assert (a[3:5]).shape == (2, 7) and len(a) == 256
    T = a.to_transformation_matrix()
    assert np.allclose(T[:, :3, :3].numpy(), G["q_matrix"], atol=1e-6) and torch.all(T[:, 3, 3] == 1)


def test_shard_ranges_partition_the_envs():
    for n in (0, 1, 7, 4096, 4099):
        for w in (1, 2, 3, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    agg = reduce_stats([ShardStats(100, 2.0, 1.0, 0.5, 300), ShardStats(100, 4.0, 3.0, 0.7, 500)])
    assert agg["solves"] == 200 and agg["elapsed_ms"] == 4.0 and agg["solves_per_s"] == 200 / 4e-3
    assert agg["err_max"] == 0.7 and agg["iterations_mean"] == 4.0


# ------------------------------------------------------------------ C ABI
def header_symbols():
    txt = open(os.path.join(ROOT, "include", "msik.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(msik_[a-z_0-9]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(built_lib):
    lib = C.CDLL(built_lib)
    syms = header_symbols()
    assert len(syms) >= 17
    for s in syms:
        assert hasattr(lib, s), f"libmsik.so does not export {s}"
    assert sorted(L.SYMBOLS) == syms, "python binding and include/msik.h disagree"
    assert L.load().msik_version() == 111


def test_struct_layouts_match_header(built_lib):
    assert C.sizeof(L.ChainDesc) == 4 + 64 + 17 * 12 * 4 + 64 + 64
    assert C.sizeof(L.StepCfg) == 7 * 4 + 64 + 12 + 12 + 4
    assert C.sizeof(L.IterCfg) == 36  # + damping, guard_pivot, mapping (MSIK_VERSION 111)
    assert C.sizeof(L.AuxCfg) == 4 + 5 * 16 * 4  # msik_aux_cfg


def _desc(robot):
    T = U.build_chain_table(robot[0], robot[1])
    d = L.ChainDesc()
    d.n_joints = T.n
    pk = T.packed_f32()
    for i in range(T.n):
        d.joint_type[i] = int(T.joint_types[i])
    for i in range(T.n + 1):
        for k in range(12):
            d.pre[i][k] = float(pk[i, k])
    return d


def test_chain_create_matches_specialisations_without_gpu(built_lib):
    lib = L.load()
    for robot, name in ((PANDA, "panda_tcp"), (FETCH, "fetch_gripper"), (XARM7, "xarm7_base")):
        h = C.c_void_p()
        d = _desc(robot)
        assert lib.msik_chain_create(C.byref(d), C.byref(h)) == 0
        buf = C.create_string_buffer(64)
        assert lib.msik_chain_kernel_name(h, buf, 64) == 0
        assert buf.value.decode() == name
        assert lib.msik_chain_destroy(h) == 0
    d = _desc(PANDA)
    d.pre[2][9] = 0.123  # a different robot -> generic table-driven kernel
    h = C.c_void_p()
    assert lib.msik_chain_create(C.byref(d), C.byref(h)) == 0
    buf = C.create_string_buffer(64)
    lib.msik_chain_kernel_name(h, buf, 64)
    assert buf.value.decode() == "generic7r"  # 7 revolute joints: the all-revolute table-driven kernels
    lib.msik_chain_destroy(h)
    d.joint_type[1] = 2  # a prismatic joint in the chain: the joint-type-dispatching ones
    assert lib.msik_chain_create(C.byref(d), C.byref(h)) == 0
    lib.msik_chain_kernel_name(h, buf, 64)
    assert buf.value.decode() == "generic7"
    lib.msik_chain_destroy(h)


def test_argument_errors_are_reported_not_crashed(built_lib):
    lib = L.load()
    h = C.c_void_p()
    d = _desc(PANDA)
    d.n_joints = 0
    assert lib.msik_chain_create(C.byref(d), C.byref(h)) == -22
    assert b"n_joints" in lib.msik_last_error()
    d = _desc(PANDA)
    d.joint_type[1] = 7
    assert lib.msik_chain_create(C.byref(d), C.byref(h)) == -22
    d = _desc(PANDA)
    assert lib.msik_chain_create(C.byref(d), C.byref(h)) == 0
    cfg = L.StepCfg()
    cfg.n_envs, cfg.qpos_stride, cfg.target_stride, cfg.action_stride = 4, 9, 9, 7
    cfg.flags = L.F_USE_DELTA
    cfg.ctrl_mask = 0x7F
    for i in range(7):
        cfg.anc_col[i] = i
    # NULL buffers
    assert lib.msik_delta_step(h, C.byref(cfg), None, None, None, None, None, None, None) == -22
    cfg.anc_col[3] = 11  # column outside the row
    assert lib.msik_delta_step(h, C.byref(cfg), 16, None, 16, 16, None, None, None) == -22
    assert b"anc_col" in lib.msik_last_error()
    cfg.anc_col[3] = 3
    cfg.ctrl_mask = 0x1FF  # more bits than joints
    assert lib.msik_delta_step(h, C.byref(cfg), 16, None, 16, 16, None, None, None) == -22
    cfg.ctrl_mask = 0x7F
    cfg.flags = L.F_USE_DELTA | L.F_USE_TARGET  # not a path-A configuration
    assert lib.msik_delta_step(h, C.byref(cfg), 16, None, 16, 16, None, None, None) == -22
    cfg.flags = L.F_USE_DELTA
    cfg.action_offset = 3  # slice [3, 9) outside stride 7
    assert lib.msik_delta_step(h, C.byref(cfg), 16, None, 16, 16, None, None, None) == -22
    cfg.action_offset = 0
    # fused sibling columns (msik_aux_cfg): collisions and out-of-row columns are refused
    aux = L.AuxCfg()
    aux.n_cols = 2
    aux.col[0], aux.col[1] = 7, 3  # column 3 is driven by the arm already
    assert lib.msik_delta_step_fused(h, C.byref(cfg), C.byref(aux), 16, None, 16, 16, None, None, None) == -22
    assert b"written twice" in lib.msik_last_error()
    aux.col[1] = 9  # outside the 9-float row
    assert lib.msik_delta_step_fused(h, C.byref(cfg), C.byref(aux), 16, None, 16, 16, None, None, None) == -22
    aux.col[1], aux.action_col[1] = 8, 7  # action column outside the 7-float action row
    assert lib.msik_delta_step_fused(h, C.byref(cfg), C.byref(aux), 16, None, 16, 16, None, None, None) == -22
    aux.action_col[1] = 6
    cfg.flags = L.F_USE_DELTA | L.F_NO_DRIVE  # sibling columns need the drive buffer
    assert lib.msik_delta_step_fused(h, C.byref(cfg), C.byref(aux), 16, None, 16, None, 16, None, None) == -22
    cfg.flags = L.F_USE_DELTA
    cfg.ctrl_mask = 0x0F  # 4 controlled joints < 6 task dims
    assert lib.msik_delta_step(h, C.byref(cfg), 16, None, 16, 16, None, None, None) == -22
    lib.msik_chain_destroy(h)


def test_no_cpu_fallback(built_lib):
    """Kinematics on a CPU articulation must raise, never silently compute on the host."""
    from maniskill_b200.kinematics import Kinematics

    class J:
        def __init__(self, name, idx, parent):
            self.name, self.parent_link = name, parent
            self.active_index = None if idx is None else torch.tensor([idx])

    class Lk:
        def __init__(self, name, joint):
            self.name, self.joint = name, joint

    root = Lk("panda_link0", J("", None, None))
    links, prev = [root], root
    for i in range(1, 8):
        l = Lk(f"panda_link{i}", J(f"panda_joint{i}", i - 1, prev))
        links.append(l)
        prev = l
    for nm in ("panda_link8", "panda_hand", "panda_hand_tcp"):
        l = Lk(nm, J(nm + "_joint", None, prev))
        links.append(l)
        prev = l

    class Art:
        device = torch.device("cpu")
        max_dof = 9

    art = Art()
    art.links = links
    art.links_map = {l.name: l for l in links}
    with pytest.raises(L.MsikError, match="no CPU fallback"):
        Kinematics(PANDA[0], "panda_hand_tcp", art, torch.arange(7, dtype=torch.int32))
    if not torch.cuda.is_available():
        with pytest.raises(L.MsikError):
            L.require_device()


def test_presets_mirror_the_reference_controller_dicts():
    """maniskill_b200.presets: same uids (in the reference's order), joint lists and bounds as the reference robots' stock
    EE controller dicts (panda.py:192-194, fetch.py:273-278, mobile_realman.py:436-443)."""
    from maniskill_b200 import presets as P
    from maniskill_b200.controllers import PDEEPosControllerConfig, PDEEPoseControllerConfig

    assert list(P.panda_pd_ee_delta_pose()) == ["arm", "gripper"]
    assert list(P.fetch_pd_ee_delta_pose()) == ["arm", "gripper", "body", "base"]
    rm = P.realman_pd_ee_delta_pos()
    assert list(rm) == ["right_arm", "right_gripper", "left_arm", "left_gripper", "body", "base"]
    assert type(rm["right_arm"]) is PDEEPosControllerConfig and rm["right_arm"].ee_link == "r_link6" and rm["left_arm"].ee_link == "l_link6"
    assert rm["right_arm"].joint_names == [f"r_joint{i}" for i in range(1, 7)] and len(rm["right_gripper"].joint_names) == 6
    assert rm["right_gripper"].lower == -10.0 and rm["right_gripper"].upper == 10.0 and rm["body"].use_delta
    assert type(P.realman_pd_ee_delta_pose()["left_arm"]) is PDEEPoseControllerConfig
    for d in (P.panda_pd_ee_delta_pose(), P.fetch_pd_ee_delta_pose(), rm):
        for c in d.values():
            if hasattr(c, "urdf_path"):
                assert os.path.exists(c.urdf_path)
