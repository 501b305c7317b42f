"""Same public names as mani_skill/agents/controllers/__init__.py:8-13 for the EE path and the controllers that share
a CombinedController with it."""
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
from .pd_joint_vel import PDJointVelController, PDJointVelControllerConfig  # noqa: F401
from .pd_base_vel import (  # noqa: F401
    PDBaseForwardVelController,
    PDBaseForwardVelControllerConfig,
    PDBaseVelController,
    PDBaseVelControllerConfig,
)
