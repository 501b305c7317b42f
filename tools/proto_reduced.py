#!/usr/bin/env python3
"""Numerical prototype of the mid-chain-frame reduced min-norm solve ("Reduced solve").

pinv(J) a for full row rank J is the minimum-norm solution of J dq = a.  The solution set, hence its
minimum-norm element, is unchanged by any invertible row operation X: (X J) dq = X a.  Expressing the
Jacobian in the frame of a mid-chain joint m with the reference point on that joint's axis makes the columns
of joints m and m+1 CONSTANT (e6 and a 2-3 entry vector); two exact row operations turn them into unit
vectors, which splits the problem into a 4 x (c-2) block A and a 2 x (c-2) block B:

    min |u|^2 + |v|^2   s.t.  A u = a1,  B u + v = a2

solved by a 5x4 Householder QR (particular solution + null vector) and a scalar minimisation.

This script checks, in emulated fp32, that the reduced solve is as accurate as the dense 7x6 Householder
solve against the fp64 pinv.  Test infrastructure only.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
from maniskill_b200 import urdf as U  # noqa: E402

f32 = np.float32


def rz(q, dt):
    c, s = np.cos(q).astype(dt), np.sin(q).astype(dt)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=dt)


def jac_base(T, q, dt=np.float64):
    """geometric Jacobian, base frame, EE reference point (rows lin; ang)."""
    R = np.eye(3, dtype=dt); p = np.zeros(3, dtype=dt)
    zs, os_ = [], []
    for i in range(T.n):
        p = p + R @ T.t[i].astype(dt)
        R = R @ T.M[i].astype(dt)
        zs.append(R[:, 2].copy()); os_.append(p.copy())
        if T.joint_types[i] == U.JOINT_REVOLUTE:
            R = R @ rz(q[i], dt)
        else:
            p = p + R[:, 2] * dt(q[i])
    pe = p + R @ T.t_tool.astype(dt)
    J = np.zeros((6, T.n), dtype=dt)
    for i in range(T.n):
        if T.joint_types[i] == U.JOINT_REVOLUTE:
            J[:3, i] = np.cross(zs[i], pe - os_[i]); J[3:, i] = zs[i]
        else:
            J[:3, i] = zs[i]
    return J


def jac_mid(T, q, m, dt):
    """J' columns in the frame after joint m's rotation, reference point = that joint's origin;
    returns (Jm (6,n), Rb (3,3) base->F_m rotation, pe (3,) EE position in F_m)."""
    n = T.n
    z = [None] * n; o = [None] * n
    # forward
    R = np.eye(3, dtype=dt); p = np.zeros(3, dtype=dt)
    for j in range(m + 1, n):
        p = p + R @ T.t[j].astype(dt)
        R = R @ T.M[j].astype(dt)
        z[j] = R[:, 2].copy(); o[j] = p.copy()
        if T.joint_types[j] == U.JOINT_REVOLUTE:
            R = R @ rz(q[j], dt)
        else:
            p = p + R[:, 2] * dt(q[j])
    pe = p + R @ T.t_tool.astype(dt)
    # backward: G = joint frame before its own motion
    R = np.eye(3, dtype=dt); p = np.zeros(3, dtype=dt)
    for j in range(m, -1, -1):
        if T.joint_types[j] == U.JOINT_REVOLUTE:
            R = R @ rz(-q[j], dt)
        else:
            p = p - R[:, 2] * dt(q[j])
        z[j] = R[:, 2].copy(); o[j] = p.copy()
        R = R @ T.M[j].astype(dt).T
        p = p - R @ T.t[j].astype(dt)
    Rb = R
    J = np.zeros((6, n), dtype=dt)
    for j in range(n):
        if T.joint_types[j] == U.JOINT_REVOLUTE:
            J[:3, j] = np.cross(o[j], z[j]); J[3:, j] = z[j]
        else:
            J[:3, j] = z[j]
    return J, Rb, pe


def householder_min_norm(A, b, dt=f32):
    """x = min-norm solution of A^T x = b, A (NR, NC); mirrors msik_math.cuh min_norm_solve."""
    A = A.astype(dt).copy(); NR, NC = A.shape
    tau = np.zeros(NC, dt); rd = np.zeros(NC, dt); rinv = np.zeros(NC, dt)
    for k in range(NC):
        sigma = dt(0)
        for r in range(k, NR):
            sigma = dt(sigma + A[r, k] * A[r, k])
        nrm = dt(np.sqrt(sigma))
        alpha = dt(-np.copysign(nrm, A[k, k]))
        v0 = dt(A[k, k] - alpha)
        den = dt(-alpha * v0)
        t = dt(1) / dt(den + dt(1e-37))
        tau[k] = t; rd[k] = alpha; rinv[k] = dt(-v0 * t)
        A[k, k] = v0
        for j in range(k + 1, NC):
            s = dt(0)
            for r in range(k, NR):
                s = dt(s + A[r, k] * A[r, j])
            s = dt(s * t)
            for r in range(k, NR):
                A[r, j] = dt(A[r, j] - s * A[r, k])
    y = np.zeros(NC, dt)
    for k in range(NC):
        acc = dt(b[k])
        for i in range(k):
            acc = dt(acc - A[i, k] * y[i])
        y[k] = dt(acc * rinv[k])
    x = np.zeros(NR, dt); x[:NC] = y

    def apply_q(x):
        for k in range(NC - 1, -1, -1):
            s = dt(0)
            for r in range(k, NR):
                s = dt(s + A[r, k] * x[r])
            s = dt(s * tau[k])
            for r in range(k, NR):
                x[r] = dt(x[r] - s * A[r, k])
        return x

    x = apply_q(x)
    null = None
    if NR == NC + 1:
        e = np.zeros(NR, dt); e[NR - 1] = 1
        null = apply_q(e)
    return x, null


def reduced_solve(T, q, a, m, ctrl, dt=f32):
    """a = (v, w) in the base frame; ctrl = list of controlled joint indices (must contain m, m+1)."""
    J, Rb, pe = jac_mid(T, q, m, dt)
    w = (Rb @ a[3:].astype(dt)).astype(dt)
    v = (Rb @ a[:3].astype(dt) + np.cross(pe, w)).astype(dt)
    rhs = np.concatenate([v, w]).astype(dt)
    # constant columns (evaluated at generation time in fp64)
    Jc, _, _ = jac_mid(T, np.zeros(T.n), m, np.float64)
    c1 = Jc[:, m + 1]
    assert np.allclose(Jc[:, m], [0, 0, 0, 0, 0, 1])
    p = int(np.argmax(np.abs(c1[:5])))
    X = np.eye(6)
    for r in range(6):
        if r != p:
            X[r, p] = -c1[r] / c1[p]
    X[p, p] = 1.0 / c1[p]
    Jr = (X.astype(dt) @ J).astype(dt); rr = (X.astype(dt) @ rhs).astype(dt)
    K = [r for r in range(5) if r != p]
    ucols = [j for j in ctrl if j not in (m, m + 1)]
    A = Jr[np.ix_(K, ucols)]; B = Jr[np.ix_([p, 5], ucols)]
    a1 = rr[K]; a2 = rr[[p, 5]]
    up, nv = householder_min_norm(A.T.copy(), a1, dt)
    wv = (a2 - B @ up).astype(dt)
    if nv is not None:
        Bn = (B @ nv).astype(dt)
        t = dt(Bn @ wv) / dt(dt(1) + Bn @ Bn)
        u = (up + nv * t).astype(dt); vv = (wv - Bn * t).astype(dt)
    else:
        u, vv = up, wv
    dq = np.zeros(T.n, dt)
    dq[ucols] = u; dq[m + 1] = vv[0]; dq[m] = vv[1]
    return dq[ctrl]


def main():
    assets = os.path.join(HERE, "..", "maniskill_b200", "assets", "robots")
    cases = [("panda_v2_kin.urdf", "panda_hand_tcp", list(range(7)), 3),
             ("fetch_kin.urdf", "gripper_link", list(range(4, 11)), 7),
             ("xarm7_ability_right_hand_kin.urdf", "base", list(range(7)), 3),
             ("realman_dual_65B_kin.urdf", "r_link6", list(range(3, 9)), 5)]
    rng = np.random.default_rng(0)
    for fixture, ee, ctrl, m in cases:
        T = U.build_chain_table(os.path.join(assets, fixture), ee)
        lo = np.where(np.isfinite(T.lower), T.lower, -np.pi); hi = np.where(np.isfinite(T.upper), T.upper, np.pi)
        lo = np.maximum(lo, -2 * np.pi); hi = np.minimum(hi, 2 * np.pi)
        e_d, e_r, conds = [], [], []
        for _ in range(400):
            q = rng.uniform(lo + 0.1 * (hi - lo), hi - 0.1 * (hi - lo))
            a = np.concatenate([rng.uniform(-0.1, 0.1, 3), rng.uniform(-0.06, 0.06, 3)])
            J = jac_base(T, q)[:, ctrl]
            ref = np.linalg.pinv(J) @ a
            cond = np.linalg.cond(J)
            if cond > 50:
                continue
            J32 = jac_base(T, q.astype(f32), f32)[:, ctrl]
            xd, _ = householder_min_norm(J32.T.copy(), a.astype(f32))
            xr = reduced_solve(T, q.astype(f32), a.astype(f32), m, ctrl)
            # sanity in fp64
            x64 = reduced_solve(T, q, a, m, ctrl, np.float64)
            assert np.abs(x64 - ref).max() < 1e-9, (np.abs(x64 - ref).max(), cond)
            e_d.append(np.abs(xd - ref).max()); e_r.append(np.abs(xr - ref).max()); conds.append(cond)
        e_d, e_r = np.array(e_d), np.array(e_r)
        print(f"{fixture:40s} m={m} n={len(e_d)} cond<=50: dense fp32 max {e_d.max():.2e} mean {e_d.mean():.2e} | "
              f"reduced fp32 max {e_r.max():.2e} mean {e_r.mean():.2e}")


if __name__ == "__main__":
    main()
