"""The reduced path-A solve (maniskill_b200/chaingen.py gen_reduced, csrc/msik_math.cuh reduced_min_norm) on the CPU.

The generated ``Chain::reduced`` members of csrc/generated/chains.cuh are plain C++ apart from the CUDA
qualifiers, so they are compiled here with g++ (qualifiers defined away) and evaluated in fp32; the small
constrained problem they set up is then solved in fp64 numpy and must reproduce the oracle's
``q0 + pinv(J) a`` (kinematics.py:156-172) on the controlled joints.  Also: the numpy prototype of the whole
derivation (tools/proto_reduced.py) against pinv in fp64, and that the committed chains.cuh is what the
generator emits today."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import FETCH, FLOATING, PANDA, REALMAN_L, REALMAN_R, ROOT, XARM7
from oracle import oracle as O

sys.path.insert(0, os.path.join(ROOT, "tools"))

CHAINS = {  # struct name -> (robot fixture, c0, nr)
    "panda_tcp": (PANDA, 0, 7),
    "fetch_gripper": (FETCH, 4, 7),
    "xarm7_base": (XARM7, 0, 7),
    "floating_gripper_tcp": (FLOATING, 0, 6),
    "realman_r_link6": (REALMAN_R, 3, 6),
    "realman_l_link6": (REALMAN_L, 3, 6),
}

_HARNESS = r"""
#include <cmath>
#include <cstdint>
#define __device__
#define __forceinline__ inline
struct msik_chain_desc {};
#include "generated/chains.cuh"
template <class Ch>
static int run(const float* q, const float* b, float* out) {
  constexpr int N = Ch::N, NU = Ch::kRedNU;
  if (!Ch::kHasReduced) return -1;
  float qq[N], sn[N], cs[N], bb[6];
  for (int i = 0; i < N; ++i) { qq[i] = q[i]; sn[i] = sinf(q[i]); cs[i] = cosf(q[i]); }
  for (int k = 0; k < 6; ++k) bb[k] = b[k];
  float At[NU][4], Bm[2][NU], a1[4], a2[2];
  Ch::reduced(qq, sn, cs, bb, At, Bm, a1, a2);
  int o = 0;
  for (int i = 0; i < NU; ++i) for (int k = 0; k < 4; ++k) out[o++] = At[i][k];
  for (int r = 0; r < 2; ++r) for (int i = 0; i < NU; ++i) out[o++] = Bm[r][i];
  for (int k = 0; k < 4; ++k) out[o++] = a1[k];
  for (int k = 0; k < 2; ++k) out[o++] = a2[k];
  out[o++] = (float)Ch::kRedSkip; out[o++] = (float)Ch::kRedZB;
  return (Ch::kRedC0 << 24) | (Ch::kRedNR << 16) | (NU << 8) | Ch::kRedM;
}
template <class Ch>
static void scat(const float* u, const float* v, float* dq) {
  float uu[Ch::kRedNU], vv[2] = {v[0], v[1]}, d[Ch::kRedNR];
  for (int i = 0; i < Ch::kRedNU; ++i) uu[i] = u[i];
  Ch::reduced_scatter(uu, vv, d);
  for (int r = 0; r < Ch::kRedNR; ++r) dq[r] = d[r];
}
template <class Ch>
static int lin(const float* q, const float* b, float* out) {
  constexpr int N = Ch::N, NR = Ch::kLinNR;
  if (!Ch::kHasLinear) return -1;
  float qq[N], sn[N], cs[N], bb[3];
  for (int i = 0; i < N; ++i) { qq[i] = q[i]; sn[i] = sinf(q[i]); cs[i] = cosf(q[i]); }
  for (int k = 0; k < 3; ++k) bb[k] = b[k];
  float At[NR][3], a3[3];
  Ch::linear(qq, sn, cs, bb, At, a3);
  int o = 0;
  for (int r = 0; r < NR; ++r) for (int k = 0; k < 3; ++k) out[o++] = At[r][k];
  for (int k = 0; k < 3; ++k) out[o++] = a3[k];
  return (Ch::kLinC0 << 16) | (NR << 8) | (Ch::kLinLastZero ? 1 : 0);
}
#define X(name, n) \
  extern "C" int lin_##name(const float* q, const float* b, float* out) { return lin<MsikChain_##name>(q, b, out); } \
  extern "C" int red_##name(const float* q, const float* b, float* out) { return run<MsikChain_##name>(q, b, out); } \
  extern "C" void scat_##name(const float* u, const float* v, float* dq) { scat<MsikChain_##name>(u, v, dq); }
MSIK_FOREACH_SPECIALIZED_CHAIN(X)
"""


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    d = tmp_path_factory.mktemp("reduced")
    src = d / "harness.cpp"
    src.write_text(_HARNESS)
    so = d / "harness.so"
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "maniskill_b200", "csrc"), str(src), "-o", str(so)],
                   check=True, capture_output=True)
    return C.CDLL(str(so))


def _solve_reduced(At, Bm, a1, a2):
    """fp64 restatement of reduced_min_norm: min |u|^2 + |a2 - B u|^2  s.t.  A u = a1."""
    A = At.T
    up = np.linalg.pinv(A) @ a1
    w = a2 - Bm @ up
    if A.shape[1] > 4:
        n = np.linalg.svd(A)[2][-1]  # unit null vector of A
        g = Bm @ n
        t = (g @ w - n @ up) / (n @ n + g @ g)
        return up + n * t, w - g * t
    return up, w


@pytest.mark.parametrize("name", list(CHAINS))
def test_generated_reduced_system_reproduces_pinv(harness, name):
    robot, c0, nr = CHAINS[name]
    fr = O.load_chain(robot[0], robot[1])
    lim = np.clip(O.joint_limits(fr), -3.0, 3.0)
    n = len(lim)
    rng = np.random.default_rng(11)
    fn = getattr(harness, "red_" + name)
    fn.restype = C.c_int
    nu = nr - 2
    worst, checked = 0.0, 0
    for _ in range(200):
        q = rng.uniform(lim[:, 0] + 0.1 * (lim[:, 1] - lim[:, 0]), lim[:, 1] - 0.1 * (lim[:, 1] - lim[:, 0]))
        a = np.concatenate([rng.uniform(-0.1, 0.1, 3), rng.uniform(-0.06, 0.06, 3)])
        mask = np.zeros(n, bool)
        mask[c0:c0 + nr] = True
        J = O.jacobian(fr, q[None])[0][0][:, mask]
        if np.linalg.cond(J) > 50:
            continue
        ref = (O.ik_delta_step(fr, q[None], a[None], col_mask=mask)[0] - q[mask])
        out = np.zeros(4 * nu + 2 * nu + 8, np.float32)
        q32, a32 = q.astype(np.float32), a.astype(np.float32)
        code = fn(q32.ctypes.data_as(C.POINTER(C.c_float)), a32.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_float)))
        assert code >= 0, "chain has no reduced solve"
        assert (code >> 24, (code >> 16) & 255, (code >> 8) & 255) == (c0, nr, nu)
        o = out.astype(np.float64)
        At = o[:4 * nu].reshape(nu, 4)
        Bm = o[4 * nu:6 * nu].reshape(2, nu)
        a1, a2 = o[6 * nu:6 * nu + 4], o[6 * nu + 4:6 * nu + 6]
        skip, zb = int(o[-2]), int(o[-1])
        assert (At[nu - 1, :skip] == 0).all()  # what reduced_min_norm<NU, SKIP, ZB> is allowed to leave out
        assert all(Bm[r, i] == 0 for r in range(2) for i in range(nu) if (zb >> (r * nu + i)) & 1)
        u, v = _solve_reduced(At, Bm, a1, a2)
        u32, v32, dq32 = u.astype(np.float32), v.astype(np.float32), np.zeros(nr, np.float32)
        fp = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
        getattr(harness, "scat_" + name)(fp(u32), fp(v32), fp(dq32))
        dq = dq32.astype(np.float64)
        worst = max(worst, np.abs(dq - ref).max())
        checked += 1
    assert checked > 50
    assert worst < 1e-5, worst  # fp32 system (generated code, sinf / cosf), fp64 solve


@pytest.mark.parametrize("name", list(CHAINS))
def test_generated_linear_system_reproduces_pinv(harness, name):
    """pd_ee_delta_pos: the frame-optimised linear Jacobian (gen_linear) gives the same min-norm step as J[:, 0:3]."""
    robot, c0, nr = CHAINS[name]
    fr = O.load_chain(robot[0], robot[1])
    lim = np.clip(O.joint_limits(fr), -3.0, 3.0)
    n = len(lim)
    rng = np.random.default_rng(12)
    fn = getattr(harness, "lin_" + name)
    fn.restype = C.c_int
    mask = np.zeros(n, bool)
    mask[c0:c0 + nr] = True
    worst = 0.0
    for _ in range(200):
        q = rng.uniform(lim[:, 0] + 0.1 * (lim[:, 1] - lim[:, 0]), lim[:, 1] - 0.1 * (lim[:, 1] - lim[:, 0]))
        a = rng.uniform(-0.1, 0.1, 3)
        ref = O.ik_delta_step(fr, q[None], a[None], pos_only=True, col_mask=mask)[0] - q[mask]
        out = np.zeros(3 * nr + 3, np.float32)
        q32, a32 = q.astype(np.float32), a.astype(np.float32)
        code = fn(q32.ctypes.data_as(C.POINTER(C.c_float)), a32.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_float)))
        assert code >= 0 and (code >> 16, (code >> 8) & 255) == (c0, nr)
        At = out[:3 * nr].astype(np.float64).reshape(nr, 3)
        a3 = out[3 * nr:].astype(np.float64)
        if code & 1:
            assert (At[-1] == 0).all()  # what min_norm_solve<NR, 3, SKIP = 3> leaves out
        assert abs(np.linalg.norm(a3) - np.linalg.norm(a)) < 1e-6  # the task frame is only rotated
        dq = np.linalg.pinv(At.T) @ a3
        worst = max(worst, np.abs(dq - ref).max())
    assert worst < 5e-6, worst


def test_prototype_matches_pinv_fp64():
    import proto_reduced as P
    from maniskill_b200 import urdf as U

    T = U.build_chain_table(PANDA[0], PANDA[1])
    rng = np.random.default_rng(3)
    for m in range(6):
        for _ in range(10):
            q = rng.uniform(-2, 2, 7)
            a = rng.uniform(-0.1, 0.1, 6)
            J = P.jac_base(T, q)
            if np.linalg.cond(J) > 200:
                continue
            ref = np.linalg.pinv(J) @ a
            got = P.reduced_solve(T, q, a, m, list(range(7)), np.float64)
            assert np.abs(got - ref).max() < 1e-8, (m, np.abs(got - ref).max())


def test_committed_chains_header_is_current():
    import gen_chain_kernels as G
    from maniskill_b200 import urdf as U
    from maniskill_b200.chaingen import gen_chain

    text = open(G.OUT).read()
    for name, fixture, ee, ctrl in G.CHAINS:
        T = U.build_chain_table(os.path.join(G.ASSETS, fixture), ee)
        assert gen_chain(name, T, ctrl) in text, f"{name}: run tools/gen_chain_kernels.py"
