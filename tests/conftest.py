import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ASSETS = os.path.join(ROOT, "maniskill_b200", "assets", "robots")

PANDA = (os.path.join(ASSETS, "panda_v2_kin.urdf"), "panda_hand_tcp", [f"panda_joint{i}" for i in range(1, 8)])
FETCH = (
    os.path.join(ASSETS, "fetch_kin.urdf"),
    "gripper_link",
    ["shoulder_pan_joint", "shoulder_lift_joint", "upperarm_roll_joint", "elbow_flex_joint", "forearm_roll_joint",
     "wrist_flex_joint", "wrist_roll_joint"],
)
# the fork's dual-arm mobile robot (agents/robots/realman/mobile_realman.py:137-138): arm joints only are controlled
REALMAN_R = (os.path.join(ASSETS, "realman_dual_65B_kin.urdf"), "r_link6", [f"r_joint{i}" for i in range(1, 7)])
REALMAN_L = (os.path.join(ASSETS, "realman_dual_65B_kin.urdf"), "l_link6", [f"l_joint{i}" for i in range(1, 7)])
# FloatingPandaGripper (agents/robots/floating_panda_gripper/floating_panda_gripper.py:41-48): six root joints, all controlled
FLOATING = (os.path.join(ASSETS, "panda_v2_gripper_kin.urdf"), "panda_hand_tcp",
            ["root_x_axis_joint", "root_y_axis_joint", "root_z_axis_joint", "root_x_rot_joint", "root_y_rot_joint", "root_z_rot_joint"])
XARM7 = (os.path.join(ASSETS, "xarm7_ability_right_hand_kin.urdf"), "base", [f"joint{i}" for i in range(1, 8)])


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def built_lib():
    """libmsik.so, built in-tree if stale (nvcc cross-compiles without a GPU)."""
    from maniskill_b200 import build

    return build.build()
