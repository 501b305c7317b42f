"""The C ABI from a plain C99 host (examples/c_host/msik_host_example.c): no Python, no torch on the data path."""
import os
import shutil
import subprocess

import pytest

from conftest import ROOT

SRC = os.path.join(ROOT, "examples", "c_host", "msik_host_example.c")
FIX = os.path.join(ROOT, "tests", "golden", "c_host_panda.bin")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _build(tmp_path, built_lib):
    exe = str(tmp_path / "msik_host_example")
    libdir = os.path.dirname(built_lib)
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(CUDA, "include"), SRC,
           "-L" + libdir, "-lmsik", "-L" + os.path.join(CUDA, "lib64"), "-lcudart", "-lm", "-Wl,-rpath," + libdir, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_header_is_c99_and_the_example_links(tmp_path, built_lib):
    """include/msik.h compiles as C (not only C++), the example links against libmsik.so, and without a GPU the
    product path fails loudly (MSIK_ENODEV) instead of falling back to anything."""
    import torch

    exe = _build(tmp_path, built_lib)
    if not torch.cuda.is_available():
        r = subprocess.run([exe, FIX], capture_output=True, text=True)
        assert r.returncode == 2 and "msik error -19" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_c_host_example_matches_the_oracle_fixture(tmp_path, built_lib):
    exe = _build(tmp_path, built_lib)
    r = subprocess.run([exe, FIX], capture_output=True, text=True)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, (r.stdout, r.stderr)
    assert "kernel for this chain: panda_tcp" in r.stdout
