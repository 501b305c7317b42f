"""Same public names as mani_skill/agents/controllers/__init__.py:8-13 for the EE path."""
from .base_controller import BaseController, CombinedController, ControllerConfig, DictController  # noqa: F401
from .pd_ee_pose import (  # noqa: F401
    PDEEPosController,
    PDEEPosControllerConfig,
    PDEEPoseController,
    PDEEPoseControllerConfig,
)
from .pd_joint_pos import (  # noqa: F401
    PDJointPosController,
    PDJointPosControllerConfig,
    PDJointPosMimicController,
    PDJointPosMimicControllerConfig,
)
