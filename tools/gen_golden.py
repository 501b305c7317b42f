#!/usr/bin/env python3
"""Generate golden vectors by IMPORTING the reference's own pure-torch module
mani_skill/utils/geometry/rotation_conversions.py (it imports standalone: torch only).

Runs only in the build container (/root/reference is not on the GPU box); the
output tests/golden/rotation_golden.npz is committed together with this script.
Also stores the known-answer FK / pinv vectors of SURVEY.md Appendix A.1 (computed
there in fp64 from panda_v2.urdf / fetch.urdf and verified by finite differences).
"""
import importlib.util
import os
import sys

import numpy as np
import torch

REF = "/root/reference/mani_skill/utils/geometry/rotation_conversions.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "rotation_golden.npz")


def main() -> int:
    spec = importlib.util.spec_from_file_location("ref_rotation_conversions", REF)
    rc = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rc)
    g = torch.Generator().manual_seed(2022)
    N = 256
    euler = (torch.rand((N, 3), generator=g, dtype=torch.float64) * 2 - 1) * 3.1
    euler[:8] *= 0.03  # small deltas like the controller's 0.1 rad actions
    euler[8] = torch.tensor([0.0, 0.0, 0.0])
    euler[9] = torch.tensor([3.141592653589793, 0.0, 0.0])
    euler[10] = torch.tensor([0.0, 3.141592653589793, 0.0])
    euler[11] = torch.tensor([0.0, 0.0, 3.141592653589793])
    mat = rc.euler_angles_to_matrix(euler, "XYZ")
    quat = rc.matrix_to_quaternion(mat)
    qa = rc.random_quaternions(N, dtype=torch.float64)
    qb = rc.random_quaternions(N, dtype=torch.float64)
    pts = torch.randn((N, 3), generator=g, dtype=torch.float64)
    out = dict(
        euler=euler.numpy(),
        euler_matrix=mat.numpy(),
        euler_quat=quat.numpy(),
        qa=qa.numpy(),
        qb=qb.numpy(),
        pts=pts.numpy(),
        q_mul=rc.quaternion_multiply(qa, qb).numpy(),
        q_raw_mul=rc.quaternion_raw_multiply(qa, qb).numpy(),
        q_apply=rc.quaternion_apply(qa, pts).numpy(),
        q_axis_angle=rc.quaternion_to_axis_angle(qa).numpy(),
        q_matrix=rc.quaternion_to_matrix(qa).numpy(),
        q_from_matrix=rc.matrix_to_quaternion(rc.quaternion_to_matrix(qa)).numpy(),
        # Pose.__mul__ / Pose.inv (pose.py:187-220) composed from the reference functions
        pose_mul_q=rc.quaternion_multiply(qa, qb).numpy(),
        pose_mul_p=(pts + rc.quaternion_apply(qa, pts.flip(0))).numpy(),
        pose_inv_p=rc.quaternion_apply(qa * torch.tensor([1.0, -1, -1, -1], dtype=torch.float64), -pts).numpy(),
        # SURVEY.md Appendix A.1 known answers (Panda panda_hand_tcp)
        kat_q=np.array([[0, np.pi / 8, 0, -5 * np.pi / 8, 0, 3 * np.pi / 4, -np.pi / 4], [0.3, -0.5, 0.2, -1.8, 0.4, 1.9, -0.7]]),
        kat_p=np.array([[0.615013339, 0, 0.169781855], [0.361693915, 0.302095527, 0.709162188]]),
        kat_quat=np.array([[0, 0.707106781, 0.707106781, 0], [0.003648218, 0.520512095, 0.789844167, 0.324345560]]),
        kat_action=np.array([0.05, -0.02, 0.03, 0.01, -0.04, 0.02]),
        kat_dq=np.array(
            [
                [-0.033668614, 0.063100661, -0.000886154, 0.181088728, 0.014621719, -0.077988067, -0.064826430],
                [-0.013691742, 0.118633574, -0.058086395, 0.153461921, -0.010851930, -0.026575214, -0.082620006],
            ]
        ),
        kat_sigma_rest=np.array([1.9453, 1.9104, 0.9301, 0.3662, 0.3547, 0.1888]),
        kat_p_zero=np.array([0.088, 0.0, 0.8226]),
        kat_quat_zero=np.array([0, 0.923879533, 0.382683432, 0]),
        kat_fetch_q=np.array([0, 0, 0, 0.386, -0.370, 0.562, -1.032, 0.695, 0.955, -0.1, 2.077]),
        kat_fetch_p=np.array([0.594376379, -0.587545865, 0.572822381]),
    )
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print("wrote", os.path.relpath(OUT), os.path.getsize(OUT), "B")
    return 0


if __name__ == "__main__":
    sys.exit(main())
