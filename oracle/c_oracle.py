"""ctypes binding of oracle/ik_oracle.c (fp64 C restatement; TEST INFRASTRUCTURE and
timed CPU baseline only -- see the header of ik_oracle.c)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor
from typing import Optional, Sequence

import numpy as np

from . import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
_TYPES = {"fixed": 0, "revolute": 1, "continuous": 1, "prismatic": 2}


class Frame(C.Structure):
    _fields_ = [("type", C.c_int32), ("xyz", C.c_double * 3), ("rpy", C.c_double * 3), ("axis", C.c_double * 3)]


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "ik_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def pack(frames: Sequence[O.Frame]):
    arr = (Frame * len(frames))()
    for i, f in enumerate(frames):
        arr[i].type = _TYPES[f.type]
        for k in range(3):
            arr[i].xyz[k] = f.xyz[k]
            arr[i].rpy[k] = f.rpy[k]
            arr[i].axis[k] = f.axis[k]
    return arr


def _dp(a):
    return a.ctypes.data_as(C.c_void_p)


def _fan(fn, B, threads):
    """run fn(lo, hi) over [0, B) in `threads` contiguous chunks on host threads."""
    threads = max(1, min(int(threads), B))
    if threads == 1:
        fn(0, B)
        return
    edges = np.linspace(0, B, threads + 1).astype(int)
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda i: fn(edges[i], edges[i + 1]), range(threads)))


def _mask(mask, n):
    m = np.ones(n, np.uint8) if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
    return m


def fk_jac(frames, q):
    q = np.ascontiguousarray(q, np.float64)
    B, n = q.shape
    fr = pack(frames)
    J = np.zeros((B, 6, n)); R = np.zeros((B, 3, 3)); p = np.zeros((B, 3))
    f = lib().orc_fk_jac
    f.restype = C.c_int
    for e in range(B):
        f(fr, len(frames), _dp(q[e]), _dp(J[e]), _dp(R[e]), _dp(p[e]))
    return J, R, p


def delta_step(frames, q, a, mask=None, threads: int = 1):
    q = np.ascontiguousarray(q, np.float64); a = np.ascontiguousarray(a, np.float64)
    B, n = q.shape
    m = _mask(mask, n); c = int(m.sum())
    out = np.zeros((B, c))
    fr, nc = pack(frames), a.shape[1]
    f = lib().orc_delta_step_batch
    _fan(lambda lo, hi: f(fr, len(frames), n, _dp(q[lo:hi]), _dp(a[lo:hi]), nc, _dp(m), int(hi - lo), c, _dp(out[lo:hi])), B, threads)
    return out


def iterative(frames, q, tgt, mask=None, max_it=200, lr=0.2, reg=1e-9, pos_tol=1e-3, rot_tol=1e-2, patience=2, threads=1):
    q = np.ascontiguousarray(q, np.float64); tgt = np.ascontiguousarray(tgt, np.float64)
    B, n = q.shape
    m = _mask(mask, n); c = int(m.sum())
    out = np.zeros((B, c)); iters = np.zeros(B, np.int32)
    fr = pack(frames)
    f = lib().orc_iterative_batch
    pat = -1 if patience is None else patience
    _fan(lambda lo, hi: f(fr, len(frames), n, _dp(q[lo:hi]), _dp(tgt[lo:hi]), _dp(m), int(hi - lo), c, max_it, C.c_double(lr),
                          C.c_double(reg), C.c_double(pos_tol), C.c_double(rot_tol), pat, _dp(out[lo:hi]), _dp(iters[lo:hi])),
         B, threads)
    return out, iters


def clik(frames, q, tgt, mask=None, max_it=100, eps=1e-4, dt=0.1, damp=1e-6, threads=1):
    q = np.ascontiguousarray(q, np.float64); tgt = np.ascontiguousarray(tgt, np.float64)
    B, n = q.shape
    m = _mask(mask, n)
    out = np.zeros((B, n)); ok = np.zeros(B, np.uint8)
    fr = pack(frames)
    f = lib().orc_clik_batch
    _fan(lambda lo, hi: f(fr, len(frames), n, _dp(q[lo:hi]), _dp(tgt[lo:hi]), _dp(m), int(hi - lo), max_it, C.c_double(eps),
                          C.c_double(dt), C.c_double(damp), _dp(out[lo:hi]), _dp(ok[lo:hi])), B, threads)
    return out, ok.astype(bool)
