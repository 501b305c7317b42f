"""msik_delta_step_host_noalloc: the host-buffer step with its device staging in a static __device__ array
(include/msik.h; no counterpart in the reference, added for callers that
cannot call cudaMalloc).  Called through the C ABI and compared with the existing entry point msik_delta_step_host on
two calls in a row with different inputs (the array keeps the previous call's contents), plus the size limit.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import PANDA
from helpers import PANDA_REST, make_art

pytestmark = pytest.mark.gpu
Q_TOL = 1e-5  # the joint-target tolerance of the other GPU parity tests


def _kin(B):
    from maniskill_b200.kinematics import Kinematics

    art = make_art(PANDA, B)
    act = [j.name for j in art.active_joints]
    idx = torch.tensor([act.index(n) for n in PANDA[2]], dtype=torch.int32, device="cuda")
    return Kinematics(PANDA[0], PANDA[1], art, idx)


def _inputs(B, seed):
    g = torch.Generator().manual_seed(seed)
    rest = torch.tensor(PANDA_REST + [0.04, 0.04], dtype=torch.float32)
    q = (rest[None] + 0.02 * torch.randn((B, 9), generator=g)).pin_memory()
    a = (torch.rand((B, 7), generator=g) * 2 - 1).pin_memory()
    return q, a


def test_noalloc_host_step_matches_the_pipeline_entry_point_twice_in_a_row():
    from maniskill_b200 import _lib as L

    lib = L.load()
    norm = ((-0.1,) * 3, (0.1,) * 3, -0.1)
    flags = L.F_USE_DELTA | L.F_NO_DRIVE | L.F_NORMALIZE
    stream = torch.cuda.current_stream().cuda_stream
    free_after = []
    for call, (B, seed) in enumerate([(50000, 11), (33333, 12)]):  # the second call finds the first one's data in the array
        kin = _kin(B)
        h_q, h_a = _inputs(B, seed)
        want = kin.delta_step_host(h_q, h_a, normalize=norm)  # existing entry point: msik_delta_step_host
        cfg = kin.step_cfg(B, 9, 9, 7, 0, flags, pos_low=norm[0], pos_high=norm[1], rot_scale=norm[2])
        assert 0 < lib.msik_noalloc_bytes_needed(C.byref(cfg)) <= L.NOALLOC_MAX_BYTES
        got = torch.full((B, 7), float("nan"), dtype=torch.float32).pin_memory()
        L.check(lib.msik_delta_step_host_noalloc(kin._chain, C.byref(cfg), h_q.data_ptr(), h_a.data_ptr(), got.data_ptr(), stream))
        assert torch.isfinite(got).all()
        diff = (got - want).abs().max().item()
        assert diff <= Q_TOL, (call, diff)
        # a repeat on the same inputs: identical, and the device's free memory does not move (nothing is allocated)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        again = torch.empty((B, 7), dtype=torch.float32).pin_memory()
        free0 = torch.cuda.mem_get_info()[0]
        L.check(lib.msik_delta_step_host_noalloc(kin._chain, C.byref(cfg), h_q.data_ptr(), h_a.data_ptr(), again.data_ptr(), stream))
        assert torch.cuda.mem_get_info()[0] == free0
        assert torch.equal(again, got)
        free_after.append(free0)
        print(f"noalloc call {call}: B={B} max |noalloc - pipeline| = {diff:.1e}")


def test_noalloc_rejects_batches_beyond_the_static_scratch():
    from maniskill_b200 import _lib as L

    lib = L.load()
    B = 200000  # 92 B per env -> 18.4 MB > MSIK_NOALLOC_MAX_BYTES (8 MiB)
    kin = _kin(8)
    cfg = kin.step_cfg(B, 9, 9, 7, 0, L.F_USE_DELTA | L.F_NO_DRIVE)
    assert lib.msik_noalloc_bytes_needed(C.byref(cfg)) > L.NOALLOC_MAX_BYTES
    buf = np.zeros(16, np.float32)
    rc = lib.msik_delta_step_host_noalloc(kin._chain, C.byref(cfg), buf.ctypes.data, buf.ctypes.data, buf.ctypes.data, None)
    assert rc == -22  # MSIK_EINVAL, before any buffer is touched
    msg = lib.msik_last_error().decode()
    assert "MSIK_NOALLOC_MAX_BYTES" in msg and str(B) in msg
