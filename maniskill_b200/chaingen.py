"""Symbolic generator of chain-specialised FK + geometric-Jacobian device code.

Shared by tools/gen_chain_kernels.py (the robots compiled into libmsik.so, csrc/generated/chains.cuh) and
maniskill_b200/jit.py (a specialisation compiled at run time for any other URDF chain).  A tiny symbolic
evaluator tracks which scalars are compile-time constants, folds 0 / +-1, and emits one ``float`` statement
per live runtime scalar; nvcc contracts the mul/add chains into FFMA.
"""
from __future__ import annotations

import numpy as np

from maniskill_b200 import urdf as U


class Gen:
    """Symbolic scalars: ('c', value) constants or ('v', name, sign) runtime values."""

    def __init__(self):
        self.lines = []
        self.k = 0
        self.ops = 0

    @staticmethod
    def c(x):
        return ("c", float(x))

    @staticmethod
    def v(name, sign=1.0):
        return ("v", name, sign)

    def fresh(self):
        self.k += 1
        return f"t{self.k}"

    @staticmethod
    def lit(x):
        return f"{np.float32(x)!r}".replace("np.float32(", "").replace(")", "") + "f"

    def sumprod(self, terms):
        """sum over terms of product of factors -> symbolic scalar."""
        const = 0.0
        live = []  # (coef, [names])
        for fac in terms:
            coef, names = 1.0, []
            for f in fac:
                if f[0] == "c":
                    coef *= f[1]
                else:
                    coef *= f[2]
                    names.append(f[1])
            if coef == 0.0:
                continue
            if not names:
                const += coef
            else:
                live.append((coef, names))
        if not live:
            return self.c(const)
        if len(live) == 1 and const == 0.0 and len(live[0][1]) == 1 and abs(live[0][0]) == 1.0:
            return self.v(live[0][1][0], live[0][0])
        parts = []
        for i, (coef, names) in enumerate(live):
            prod = "*".join(names)
            mag = abs(coef)
            body = prod if mag == 1.0 else f"{self.lit(mag)}*{prod}"
            self.ops += len(names) - (1 if mag == 1.0 else 0) + (1 if i > 0 else 0)
            if i == 0:
                parts.append(("-" if coef < 0 else "") + body)
            else:
                parts.append((" - " if coef < 0 else " + ") + body)
        if const != 0.0:
            parts.append((" - " if const < 0 else " + ") + self.lit(abs(const)))
            self.ops += 1
        name = self.fresh()
        self.lines.append(f"const float {name} = {''.join(parts)};")
        return self.v(name)

    def expr(self, s):
        if s[0] == "c":
            return self.lit(s[1])
        return ("-" if s[2] < 0 else "") + s[1]


def gen_chain(name: str, T: U.ChainTable) -> str:
    g = Gen()
    c = g.c
    n = T.n
    R = [[c(1.0 if i == j else 0.0) for j in range(3)] for i in range(3)]
    # Positions are tracked as  base + rel:  `base` collects every translation that comes
    # BEFORE the first revolute joint (Fetch: the +-20 m x / y base prismatics), `rel` the
    # rest.  The Jacobian lever arms pe - o_i then never see the base offset, so there is no
    # fp32 cancellation when the robot stands far from the origin (pk computes the arms in
    # the EE frame for the same effect).
    base = [c(0.0)] * 3
    p = [c(0.0)] * 3
    seen_rev = False
    zs, os_, types = [], [], []
    for i in range(n):
        M, t = T.M[i], T.t[i]
        if seen_rev:
            p = [g.sumprod([[p[r]]] + [[R[r][k], c(t[k])] for k in range(3)]) for r in range(3)]
        else:
            base = [g.sumprod([[base[r]]] + [[R[r][k], c(t[k])] for k in range(3)]) for r in range(3)]
        R = [[g.sumprod([[R[r][k], c(M[k][j])] for k in range(3)]) for j in range(3)] for r in range(3)]
        zs.append([R[r][2] for r in range(3)])
        os_.append(list(p))
        types.append(int(T.joint_types[i]))
        if T.joint_types[i] == U.JOINT_REVOLUTE:
            seen_rev = True
            ci, si = g.v(f"cs[{i}]"), g.v(f"sn[{i}]")
            nsi = g.v(f"sn[{i}]", -1.0)
            X = [R[r][0] for r in range(3)]
            Y = [R[r][1] for r in range(3)]
            for r in range(3):
                R[r][0] = g.sumprod([[ci, X[r]], [si, Y[r]]])
                R[r][1] = g.sumprod([[nsi, X[r]], [ci, Y[r]]])
        else:
            qi = g.v(f"q[{i}]")
            if seen_rev:
                p = [g.sumprod([[p[r]], [zs[i][r], qi]]) for r in range(3)]
            else:
                base = [g.sumprod([[base[r]], [zs[i][r], qi]]) for r in range(3)]
    Mt, tt = T.M_tool, T.t_tool
    pe_rel = [g.sumprod([[p[r]]] + [[R[r][k], c(tt[k])] for k in range(3)]) for r in range(3)]
    pe = [g.sumprod([[pe_rel[r]], [base[r]]]) for r in range(3)]
    Re = [[g.sumprod([[R[r][k], c(Mt[k][j])] for k in range(3)]) for j in range(3)] for r in range(3)]
    body = list(g.lines)
    g.lines = []
    jac = []
    for i in range(n):
        z = zs[i]
        axial_tool = i == n - 1 and abs(tt[0]) < 1e-12 and abs(tt[1]) < 1e-12
        if types[i] == U.JOINT_REVOLUTE and axial_tool:
            # last joint, tool offset along its own axis (Panda: hand + TCP sit on joint 7's z):
            # the lever arm is parallel to the axis, the linear column is exactly zero
            lin = [c(0.0)] * 3
            ang = z
        elif types[i] == U.JOINT_REVOLUTE:
            d = [g.sumprod([[pe_rel[r]], [os_[i][r], c(-1.0)]]) for r in range(3)]
            neg = lambda s: ("c", -s[1]) if s[0] == "c" else ("v", s[1], -s[2])
            lin = [
                g.sumprod([[z[1], d[2]], [neg(z[2]), d[1]]]),
                g.sumprod([[z[2], d[0]], [neg(z[0]), d[2]]]),
                g.sumprod([[z[0], d[1]], [neg(z[1]), d[0]]]),
            ]
            ang = z
        else:
            lin = z
            ang = [c(0.0)] * 3
        jac.append((lin, ang))
    body += g.lines
    out = []
    out.append(f"// ---- {name}: {T.n} joints, end link {T.end_link!r}; ~{g.ops} fp32 ops for FK + J ----")
    out.append(f"struct MsikChain_{name} {{")
    out.append(f"  static constexpr int N = {n};")
    out.append(f"  static constexpr bool kGeneric = false;")
    last_lin_zero = types[n - 1] == U.JOINT_REVOLUTE and abs(tt[0]) < 1e-12 and abs(tt[1]) < 1e-12
    out.append(f"  static constexpr bool kLastLinZero = {'true' if last_lin_zero else 'false'};  // J[0:3, N-1] == 0 exactly")
    out.append(f"  static constexpr uint32_t kRevoluteMask = 0x{sum(1 << i for i in range(n) if types[i] == U.JOINT_REVOLUTE):x}u;")
    out.append("  static __device__ __forceinline__ void fkjac(const msik_chain_desc&, const float (&q)[N], const float (&sn)[N],")
    out.append("      const float (&cs)[N], float (&Jl)[N][3], float (&Ja)[N][3], float (&pe)[3], float (&Re)[9]) {")
    out.append("    (void)q; (void)sn; (void)cs;")
    for ln in body:
        out.append("    " + ln)
    for i, (lin, ang) in enumerate(jac):
        for r in range(3):
            out.append(f"    Jl[{i}][{r}] = {g.expr(lin[r])}; Ja[{i}][{r}] = {g.expr(ang[r])};")
    for r in range(3):
        out.append(f"    pe[{r}] = {g.expr(pe[r])};")
    for r in range(3):
        for j in range(3):
            out.append(f"    Re[{3 * r + j}] = {g.expr(Re[r][j])};")
    out.append("  }")
    out.append("};")
    # host-side table for matching
    packed = T.packed_f32()
    out.append(f"static const int kMsikTypes_{name}[{n}] = {{{', '.join(str(t) for t in types)}}};")
    out.append(f"static const float kMsikPre_{name}[{n + 1}][12] = {{")
    for row in packed:
        out.append("  {" + ", ".join(Gen.lit(x) for x in row) + "},")
    out.append("};")
    return "\n".join(out)
