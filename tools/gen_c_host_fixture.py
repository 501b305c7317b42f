#!/usr/bin/env python3
"""Fixture of examples/c_host/msik_host_example.c: chain description, step configuration, inputs in the PhysX
layout and the fp64 oracle's joint targets for 256 near-rest Panda envs (tests/golden/c_host_panda.bin).
CPU only: URDF table (maniskill_b200/urdf.py) + oracle/oracle.py."""
import ctypes as C
import os
import struct
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from maniskill_b200 import _lib as L  # noqa: E402  (ctypes mirrors of the header structs; the library is not loaded)
from maniskill_b200.urdf import build_chain_table  # noqa: E402
from oracle import oracle as O  # noqa: E402

URDF = os.path.join(ROOT, "maniskill_b200", "assets", "robots", "panda_v2_kin.urdf")
EE = "panda_hand_tcp"
REST = [0.0, np.pi / 8, 0, -np.pi * 5 / 8, 0, np.pi * 3 / 4, np.pi / 4]  # agents/robots/panda/panda.py:35-48


def main():
    B, qs, astride = 256, 9, 7
    tab = build_chain_table(URDF, EE)
    n = tab.n
    desc = L.ChainDesc()
    desc.n_joints = n
    packed = tab.packed_f32()
    for i in range(n):
        desc.joint_type[i] = int(tab.joint_types[i])
        desc.lower[i] = float(np.float32(tab.lower[i]))
        desc.upper[i] = float(np.float32(tab.upper[i]))
    for i in range(n + 1):
        for k in range(12):
            desc.pre[i][k] = float(packed[i, k])
    cfg = L.StepCfg()
    cfg.n_envs, cfg.qpos_stride, cfg.target_stride, cfg.action_stride, cfg.action_offset = B, qs, qs, astride, 0
    cfg.flags = L.F_USE_DELTA | L.F_NORMALIZE | L.F_NO_DRIVE
    cfg.ctrl_mask = (1 << n) - 1
    for i in range(n):
        cfg.anc_col[i] = i
    for k in range(3):
        cfg.pos_low[k], cfg.pos_high[k] = -0.1, 0.1
    cfg.rot_scale = -0.1
    rng = np.random.default_rng(2022)
    q = np.zeros((B, qs), np.float32)
    q[:, :7] = (np.asarray(REST)[None] + rng.normal(0, 0.02, (B, 7))).astype(np.float32)
    q[:, 7:] = 0.04
    a = rng.uniform(-1, 1, (B, astride)).astype(np.float32)
    fr = O.load_chain(URDF, EE)
    proc = O.pose_clip_and_scale(a[:, :6].astype(np.float64), -0.1, 0.1, -0.1)
    want = O.ik_delta_step(fr, q[:, :7].astype(np.float64), proc).astype(np.float32)
    out = os.path.join(ROOT, "tests", "golden", "c_host_panda.bin")
    with open(out, "wb") as f:
        f.write(struct.pack("<Iiiii", 0x4B49534D, B, qs, astride, n))
        f.write(bytes(desc))
        f.write(bytes(cfg))
        f.write(q.tobytes())
        f.write(a.tobytes())
        f.write(want.tobytes())
    print(f"wrote {out} ({os.path.getsize(out)} B): sizeof(msik_chain_desc)={C.sizeof(desc)} sizeof(msik_step_cfg)={C.sizeof(cfg)}")


if __name__ == "__main__":
    main()
