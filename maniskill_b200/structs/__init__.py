from .pose import Pose, vectorize_pose  # noqa: F401
