"""In-tree build of libmsik.so (hand-written sm_100a CUDA + the C ABI of include/msik.h).

One nvcc invocation per translation unit, run in parallel; objects and the shared
library land in maniskill_b200/_build/ and maniskill_b200/libmsik.so (git-ignored,
but shipped to the GPU box by gpurun).  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List, Tuple

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
# MSIK_VARIANT=name builds an experiment (with MSIK_EXTRA_NVCC_FLAGS) beside the product library:
# maniskill_b200/_build/variants/<name>/libmsik.so, selected at run time with MSIK_LIB=<path>
_VARIANT = os.environ.get("MSIK_VARIANT", "")
BUILD = os.path.join(PKG, "_build", "variants", _VARIANT) if _VARIANT else os.path.join(PKG, "_build")
LIB = os.path.join(BUILD, "libmsik.so") if _VARIANT else os.path.join(PKG, "libmsik.so")

SPECIALISED = ["panda_tcp", "fetch_gripper", "xarm7_base", "panda_stick_tcp", "floating_gripper_tcp", "realman_r_link6", "realman_l_link6"]
GENERIC = [7, 8, 12, 16]
GENERIC_REV = [6, 7]  # all-revolute chains of exactly this many joints (csrc/msik_kernels.cuh ChainGenericRev)

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-O2", "-Xptxas", "-v"]


def _sources_digest() -> str:
    h = hashlib.sha256()
    paths = []
    for root, _, files in os.walk(CSRC):
        for f in files:
            paths.append(os.path.join(root, f))
    paths.append(os.path.join(PKG, "..", "include", "msik.h"))
    paths.append(os.path.abspath(__file__))
    for p in sorted(paths):
        with open(p, "rb") as fh:
            h.update(os.path.relpath(p, PKG).encode())  # location-independent: the snapshot on the GPU box lives elsewhere
            h.update(fh.read())
    h.update(" ".join(CFLAGS + ARCH).encode())
    h.update(os.environ.get("MSIK_EXTRA_NVCC_FLAGS", "").encode())
    return h.hexdigest()


def _units() -> List[Tuple[str, str, List[str]]]:
    tu = os.path.join(CSRC, "msik_chain_tu.cu")
    units = [("msik.o", os.path.join(CSRC, "msik.cu"), [])]
    for name in SPECIALISED:
        units.append((f"chain_{name}.o", tu, [f"-DMSIK_TU_CHAIN={name}"]))
    for n in GENERIC:
        units.append((f"chain_generic{n}.o", tu, [f"-DMSIK_TU_GENERIC={n}"]))
    for n in GENERIC_REV:
        units.append((f"chain_generic{n}r.o", tu, [f"-DMSIK_TU_GENERIC_REV={n}"]))
    return units


class build_lock:
    """Exclusive inter-process lock on the build directory: bench.py runs one process per GPU, and ranks that
    find a stale library (or a missing JIT plugin) at the same moment must not compile / link into the same files."""

    def __init__(self, directory: str):
        os.makedirs(directory, exist_ok=True)
        self.path = os.path.join(directory, ".build.lock")

    def __enter__(self):
        import fcntl

        self.fh = open(self.path, "w")
        fcntl.flock(self.fh, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl

        fcntl.flock(self.fh, fcntl.LOCK_UN)
        self.fh.close()
        return False


def _fresh(digest: str) -> bool:
    stamp = os.path.join(BUILD, "digest.txt")
    return os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == digest


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile (if stale) and return the path of libmsik.so."""
    os.makedirs(BUILD, exist_ok=True)
    digest = _sources_digest()
    if not force and _fresh(digest):
        return LIB
    with build_lock(BUILD):
        if not force and _fresh(digest):  # another process built it while we waited for the lock
            return LIB
        return _build_locked(digest, verbose)


def _build_locked(digest: str, verbose: bool) -> str:
    stamp = os.path.join(BUILD, "digest.txt")
    extra = os.environ.get("MSIK_EXTRA_NVCC_FLAGS", "").split()

    def compile_one(unit):
        obj, src, defs = unit
        out = os.path.join(BUILD, obj)
        cmd = [NVCC, *ARCH, *CFLAGS, *extra, *defs, "-c", src, "-o", out]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(BUILD, obj + ".log")
        with open(log, "w") as fh:
            fh.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {obj}:\n{r.stdout}\n{r.stderr}")
        return out

    units = _units()
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, units))
    tmp = LIB + f".{os.getpid()}.tmp"
    cmd = [NVCC, *ARCH, "-shared", "-o", tmp, *objs, "-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB)  # readers that dlopen'ed the old file keep their inode
    with open(stamp, "w") as fh:
        fh.write(digest)
    if verbose:
        print(f"built {LIB}")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
