"""ctypes binding of libmsik.so (include/msik.h).

The CUDA library is the product: there is no Python / CPU fallback.  Importing this
module never needs a GPU (so `-m "not gpu"` tests can check the exported symbols),
but every compute entry point raises ``MsikError`` when the library, a CUDA device
of compute capability 10.x, or a kernel launch is missing.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

MSIK_MAX_JOINTS = 16

# flags (include/msik.h)
F_NORMALIZE = 0x1
F_POS_ONLY = 0x2
F_USE_DELTA = 0x4
F_USE_TARGET = 0x8
F_BODY_TRANS = 0x10
F_BODY_ROT = 0x20
F_NO_DRIVE = 0x40
F_TARGET_GIVEN = 0x80

JOINT_REVOLUTE = 1
JOINT_PRISMATIC = 2


class MsikError(RuntimeError):
    pass


class ChainDesc(C.Structure):
    _fields_ = [
        ("n_joints", C.c_int32),
        ("joint_type", C.c_int32 * MSIK_MAX_JOINTS),
        ("pre", (C.c_float * 12) * (MSIK_MAX_JOINTS + 1)),
        ("lower", C.c_float * MSIK_MAX_JOINTS),
        ("upper", C.c_float * MSIK_MAX_JOINTS),
    ]


class StepCfg(C.Structure):
    _fields_ = [
        ("n_envs", C.c_int32),
        ("qpos_stride", C.c_int32),
        ("target_stride", C.c_int32),
        ("action_stride", C.c_int32),
        ("action_offset", C.c_int32),
        ("flags", C.c_int32),
        ("ctrl_mask", C.c_uint32),
        ("anc_col", C.c_int32 * MSIK_MAX_JOINTS),
        ("pos_low", C.c_float * 3),
        ("pos_high", C.c_float * 3),
        ("rot_scale", C.c_float),
    ]


NOALLOC_MAX_BYTES = 8 << 20  # MSIK_NOALLOC_MAX_BYTES
MSIK_MAX_AUX = 16
AUX_NORMALIZE = 0x1
AUX_DELTA = 0x2


class AuxCfg(C.Structure):
    _fields_ = [
        ("n_cols", C.c_int32),
        ("col", C.c_int32 * MSIK_MAX_AUX),
        ("action_col", C.c_int32 * MSIK_MAX_AUX),
        ("flags", C.c_int32 * MSIK_MAX_AUX),
        ("mid", C.c_float * MSIK_MAX_AUX),
        ("half", C.c_float * MSIK_MAX_AUX),
    ]


VEL_JOINT, VEL_BASE, VEL_FORWARD = 0, 1, 2


class VelCfg(C.Structure):
    _fields_ = [
        ("n_envs", C.c_int32),
        ("n_cols", C.c_int32),
        ("col", C.c_int32 * MSIK_MAX_AUX),
        ("ego", C.c_int32),
        ("ori_col", C.c_int32),
        ("flags", C.c_int32),
        ("qpos_stride", C.c_int32),
        ("action_stride", C.c_int32),
        ("action_offset", C.c_int32),
        ("qvel_stride", C.c_int32),
        ("mid", C.c_float * MSIK_MAX_AUX),
        ("half", C.c_float * MSIK_MAX_AUX),
    ]


class DualCfg(C.Structure):
    _fields_ = [("chain", C.c_void_p), ("cfg", StepCfg)]


class IterCfg(C.Structure):
    _fields_ = [
        ("max_iterations", C.c_int32),
        ("patience", C.c_int32),
        ("lr", C.c_float),
        ("pos_tol", C.c_float),
        ("rot_tol", C.c_float),
        ("rank_rtol", C.c_float),
        ("damping", C.c_float),
        ("guard_pivot", C.c_float),
        ("mapping", C.c_int32),
    ]


IK_AUTO, IK_THREAD_PER_ENV, IK_LANE_PER_JOINT = 0, 1, 2


# every symbol include/msik.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "msik_version": (C.c_int, []),
    "msik_last_error": (C.c_char_p, []),
    "msik_device_check": (C.c_int, [C.POINTER(C.c_int)] * 3),
    "msik_chain_create": (C.c_int, [C.POINTER(ChainDesc), C.POINTER(_P)]),
    "msik_chain_destroy": (C.c_int, [_P]),
    "msik_chain_kernel_name": (C.c_int, [_P, C.c_char_p, C.c_int]),
    "msik_delta_step": (C.c_int, [_P, C.POINTER(StepCfg), _P, _P, _P, _P, _P, _P, _P]),
    "msik_delta_step_fused": (C.c_int, [_P, C.POINTER(StepCfg), C.POINTER(AuxCfg), _P, _P, _P, _P, _P, _P, _P]),
    "msik_delta_step_fused_vel": (C.c_int, [_P, C.POINTER(StepCfg), C.POINTER(AuxCfg), C.POINTER(VelCfg), _P, _P, _P, _P, _P, _P, _P, _P]),
    "msik_step_create": (C.c_int, [_P, C.POINTER(StepCfg), C.POINTER(AuxCfg), C.POINTER(VelCfg), _P, _P, _P, _P, C.POINTER(_P)]),
    "msik_step_create_dual": (C.c_int, [_P, C.POINTER(StepCfg), C.POINTER(DualCfg), C.POINTER(AuxCfg), C.POINTER(VelCfg), _P, _P, _P, _P,
                                        C.POINTER(_P)]),
    "msik_delta_step_dual": (C.c_int, [_P, C.POINTER(StepCfg), C.POINTER(DualCfg), C.POINTER(AuxCfg), C.POINTER(VelCfg), _P, _P, _P, _P, _P, _P]),
    "msik_step_launch": (C.c_int, [_P, _P, _P]),
    "msik_step_destroy": (C.c_int, [_P]),
    "msik_target_ik": (C.c_int, [_P, C.POINTER(StepCfg), C.POINTER(IterCfg)] + [_P] * 11),
    "msik_register_chain_ops": (C.c_int, [_P]),
    "msik_fk": (C.c_int, [_P, C.POINTER(StepCfg), _P, _P, _P, _P, _P]),
    "msik_fk_world": (C.c_int, [_P, C.POINTER(StepCfg), _P, _P, _P, _P, _P]),
    "msik_ee_pose_at_base": (C.c_int, [_P, C.c_int32, _P, _P, C.c_int32, _P, _P]),
    "msik_target_pose": (C.c_int, [C.POINTER(StepCfg), _P, _P, _P, _P]),
    "msik_pose_to_delta_action": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float, C.c_float, C.c_int32, _P, _P,
                                            C.c_int32, _P, _P, _P]),
    "msik_interp_targets": (C.c_int, [C.POINTER(StepCfg), _P, _P, C.c_float, _P, _P, _P]),
    "msik_vel_targets": (C.c_int, [C.POINTER(VelCfg), _P, _P, _P, _P, _P]),
    "msik_reset_masked": (C.c_int, [C.POINTER(StepCfg), _P, _P, _P, _P, _P, _P, _P, _P]),
    "msik_pipeline_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.POINTER(_P)]),
    "msik_pipeline_destroy": (C.c_int, [_P]),
    "msik_delta_step_host": (C.c_int, [_P, _P, C.POINTER(StepCfg), _P, _P, _P]),
    "msik_noalloc_bytes_needed": (C.c_int64, [C.POINTER(StepCfg)]),
    "msik_delta_step_host_noalloc": (C.c_int, [_P, C.POINTER(StepCfg), _P, _P, _P, _P]),
    "msik_launch_count": (C.c_int64, []),
}

# MSIK_LIB: load another build of the same library (kernel A/B experiments); default = in-tree build
LIB_PATH = os.environ.get("MSIK_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libmsik.so")
_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """dlopen libmsik.so and bind every symbol.  Raises MsikError if it is missing --
    run ``python -c 'import __graft_entry__ as g; g.build()'`` (or maniskill_b200/build.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MsikError(
            f"{LIB_PATH} not built. This package has no CPU / PyTorch fallback: build the CUDA "
            "extension with `python maniskill_b200/build.py`."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # pragma: no cover
            raise MsikError(f"libmsik.so does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().msik_last_error()
        raise MsikError(f"libmsik error {rc}: {msg.decode() if msg else '?'}")


def require_device() -> int:
    """Returns the SM count; raises when no sm_100 device is usable (no fallback)."""
    sm, maj, mnr = C.c_int(0), C.c_int(0), C.c_int(0)
    check(load().msik_device_check(C.byref(sm), C.byref(maj), C.byref(mnr)))
    return sm.value


def ptr(t) -> Optional[int]:
    """device / host pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()
