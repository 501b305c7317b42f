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
        self.deps = {}
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
        self.deps[name] = sorted({nm for _, names in live for nm in names})
        return self.v(name)

    def live_lines(self, outputs):
        """The emitted statements that the symbolic scalars in ``outputs`` depend on (dead ones dropped)."""
        need, stack = set(), [s[1] for s in outputs if s[0] == "v"]
        while stack:
            nm = stack.pop()
            if nm in need:
                continue
            need.add(nm)
            stack.extend(self.deps.get(nm, ()))
        return [ln for ln in self.lines if ln.split()[2] in need]

    def expr(self, s):
        if s[0] == "c":
            return self.lit(s[1])
        return ("-" if s[2] < 0 else "") + s[1]


def gen_chain(name: str, T: U.ChainTable, ctrl=None) -> str:
    """``ctrl`` = (c0, nr): the contiguous range of chain joints the stock controller drives (default: all);
    the reduced path-A solve (gen_reduced) is generated for exactly that range."""
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
    c0, nr = ctrl if ctrl is not None else (0, n)
    out += gen_reduced(T, c0, nr)
    out += gen_linear(T, c0, nr)
    out.append("};")
    # host-side table for matching
    packed = T.packed_f32()
    out.append(f"static const int kMsikTypes_{name}[{n}] = {{{', '.join(str(t) for t in types)}}};")
    out.append(f"static const float kMsikPre_{name}[{n + 1}][12] = {{")
    for row in packed:
        out.append("  {" + ", ".join(Gen.lit(x) for x in row) + "},")
    out.append("};")
    return "\n".join(out)


# ----------------------------------------------------------------------------------------------------------
# Reduced path-A solve.  pinv(J) a (kinematics.py:171) for full row rank J is the minimum-norm solution of
# J dq = a, and that solution is unchanged by any invertible ROW operation X: (X J) dq = X a.  Expressed in the
# frame of a mid-chain revolute joint m (after its rotation) with the reference point at that joint's origin,
# the column of joint m is e6 and the column of joint m+1 is a constant; two constant row operations make the
# latter a unit vector e_p too.  With the unknowns split into v = (dq_{m+1}, dq_m) and u (the other c - 2
# controlled joints) the system reads   A u = a1  (4 rows),   B u + v = a2  (2 rows):
#     minimise |u|^2 + |a2 - B u|^2  subject to  A u = a1
# = a (c-2) x 4 Householder QR (particular solution + null vector) and a scalar minimisation, instead of a
# c x 6 QR; FK shrinks too (only the base -> F_m rotation and lever arms relative to joint m are needed).
# tools/proto_reduced.py is the numpy prototype (fp64: identical to pinv; fp32: no worse than the dense QR).
# ----------------------------------------------------------------------------------------------------------
def _neg(s):
    return ("c", -s[1]) if s[0] == "c" else ("v", s[1], -s[2])


def _cross(g, a, b):
    return [
        g.sumprod([[a[1], b[2]], [_neg(a[2]), b[1]]]),
        g.sumprod([[a[2], b[0]], [_neg(a[0]), b[2]]]),
        g.sumprod([[a[0], b[1]], [_neg(a[1]), b[0]]]),
    ]


def _frame_chains(g, T: U.ChainTable, m: int, extra=None):
    """Joint axes z_j and origins o_j of the whole chain, the EE position pe and the base -> frame rotation Rb, all in
    F_m = the frame of joint m after its own motion (m = -1: the base frame), as symbolic scalars of ``g``.
    ``extra`` (dict, optional) receives "Re" = end-link orientation in F_m and "pb" = base origin in F_m."""
    if extra is None:
        extra = {}
    c = g.c
    n = T.n
    rev = lambda j: T.joint_types[j] == U.JOINT_REVOLUTE
    z = [None] * n
    o = [None] * n
    R = [[c(1.0 if i == j else 0.0) for j in range(3)] for i in range(3)]
    p = [c(0.0)] * 3

    def rot_right(R, ci, si):  # R * Rz(angle) for (cos, sin) = (ci, si)
        X = [R[r][0] for r in range(3)]
        Y = [R[r][1] for r in range(3)]
        for r in range(3):
            R[r][0] = g.sumprod([[ci, X[r]], [si, Y[r]]])
            R[r][1] = g.sumprod([[_neg(si), X[r]], [ci, Y[r]]])

    # forward from F_m
    for j in range(m + 1, n):
        M, t = T.M[j], T.t[j]
        p = [g.sumprod([[p[r]]] + [[R[r][k], c(t[k])] for k in range(3)]) for r in range(3)]
        R = [[g.sumprod([[R[r][k], c(M[k][i])] for k in range(3)]) for i in range(3)] for r in range(3)]
        z[j] = [R[r][2] for r in range(3)]
        o[j] = list(p)
        if rev(j):
            rot_right(R, g.v(f"cs[{j}]"), g.v(f"sn[{j}]"))
        else:
            p = [g.sumprod([[p[r]], [z[j][r], g.v(f"q[{j}]")]]) for r in range(3)]
    pe = [g.sumprod([[p[r]]] + [[R[r][k], c(T.t_tool[k])] for k in range(3)]) for r in range(3)]
    extra["Re"] = [[g.sumprod([[R[r][k], c(T.M_tool[k][i])] for k in range(3)]) for i in range(3)] for r in range(3)]
    # backward: pose of each joint's own frame (before its motion) in F_m
    R = [[c(1.0 if i == j else 0.0) for j in range(3)] for i in range(3)]
    p = [c(0.0)] * 3
    for j in range(m, -1, -1):
        if rev(j):
            rot_right(R, g.v(f"cs[{j}]"), g.v(f"sn[{j}]", -1.0))  # Rz(-q_j)
        else:
            p = [g.sumprod([[p[r]], [R[r][2], g.v(f"q[{j}]", -1.0)]]) for r in range(3)]
        z[j] = [R[r][2] for r in range(3)]
        o[j] = list(p)
        M, t = T.M[j], T.t[j]
        R = [[g.sumprod([[R[r][k], c(M[i][k])] for k in range(3)]) for i in range(3)] for r in range(3)]  # R * M^T
        p = [g.sumprod([[p[r]]] + [[R[r][k], c(-t[k])] for k in range(3)]) for r in range(3)]
    extra["pb"] = list(p)  # origin of the base frame in F_m: a base-frame point x sits at R x + pb
    return z, o, pe, R  # R: base-frame vector -> F_m


def _reduced_system(T: U.ChainTable, c0: int, nr: int, m: int, frame_out=None):
    """Symbolic reduced system for the mid joint m.  Returns (gen, A rows, B rows, a1, a2, ucols, p) or None.
    ``frame_out`` (dict, optional) receives the frame quantities the path-B variant needs (gen_reduced)."""
    g = Gen()
    c = g.c
    n = T.n
    ctrl = list(range(c0, c0 + nr))
    rev = lambda j: T.joint_types[j] == U.JOINT_REVOLUTE
    extra = {}
    z, o, pe, R = _frame_chains(g, T, m, extra)
    Rb = R  # base-frame vector -> F_m
    if frame_out is not None:
        frame_out.update(Rb=Rb, pe=pe, Re=extra["Re"], pb=extra["pb"])
    cols = {}
    for j in ctrl:
        if rev(j):
            cols[j] = _cross(g, o[j], z[j]) + list(z[j])  # z x (0 - o) = o x z
        else:
            cols[j] = list(z[j]) + [c(0.0)] * 3
    b = [g.v(f"b[{k}]") for k in range(6)]
    w = [g.sumprod([[Rb[r][k], b[3 + k]] for k in range(3)]) for r in range(3)]
    pxw = _cross(g, pe, w)
    v = [g.sumprod([[Rb[r][k], b[k]] for k in range(3)] + [[pxw[r]]]) for r in range(3)]
    rhs = v + w
    cm, cn = cols[m], cols[m + 1]
    if any(s[0] != "c" for s in cm + cn):
        return None
    cmv = np.array([s[1] for s in cm])
    cnv = np.array([s[1] for s in cn])
    if not np.allclose(cmv, [0, 0, 0, 0, 0, 1], atol=1e-12):
        return None
    pr = int(np.argmax(np.abs(cnv[:5])))
    if abs(cnv[pr]) < 0.2:
        return None  # joint m+1 (nearly) parallel to joint m through the same point: no well-scaled pivot

    def rowop(vec):
        out = []
        for r in range(6):
            if r == pr:
                out.append(g.sumprod([[vec[pr], c(1.0 / cnv[pr])]]))
            elif abs(cnv[r]) > 1e-14:
                out.append(g.sumprod([[vec[r]], [vec[pr], c(-cnv[r] / cnv[pr])]]))
            else:
                out.append(vec[r])
        return out

    if frame_out is not None:
        frame_out.update(rowop=rowop)
    ucols = [j for j in ctrl if j not in (m, m + 1)]
    K = [r for r in range(5) if r != pr]
    red = {j: rowop(cols[j]) for j in ucols}
    rr = rowop(rhs)
    A = [[red[j][r] for j in ucols] for r in K]      # 4 x NU
    Bm = [[red[j][r] for j in ucols] for r in (pr, 5)]  # 2 x NU
    a1 = [rr[r] for r in K]
    a2 = [rr[pr], rr[5]]
    return g, A, Bm, a1, a2, ucols, pr


def gen_reduced(T: U.ChainTable, c0: int, nr: int):
    """Member lines of the chain struct for the reduced solve, or the ``kHasReduced = false`` stub."""
    stub = ["  static constexpr bool kHasReduced = false;", "  static constexpr int kRedC0 = 0, kRedNR = 1, kRedNU = 4, kRedM = 0, kRedSkip = 0;",
            "  static constexpr uint32_t kRedZB = 0u;",
            "  static __device__ __forceinline__ void reduced(const float (&)[N], const float (&)[N], const float (&)[N], const float (&)[6],",
            "      float (&)[kRedNU][4], float (&)[2][kRedNU], float (&)[4], float (&)[2]) {}",
            "  static __device__ __forceinline__ void reduced_scatter(const float (&)[kRedNU], const float (&)[2], float (&)[kRedNR]) {}",
            "  static __device__ __forceinline__ void reduced_ik_sys(const float (&)[N], const float (&)[N], const float (&)[N], const float (&)[3],",
            "      const float (&)[9], float (&)[kRedNU][4], float (&)[2][kRedNU], float (&)[3], float (&)[9], float (&)[9], float (&)[3]) {}",
            "  static __device__ __forceinline__ void reduced_ik_rhs(const float (&)[3], const float (&)[3], const float (&)[3], float (&)[4], float (&)[2]) {}"]
    nu = nr - 2
    if nu not in (4, 5):
        return stub
    best = None
    for m in range(c0, c0 + nr - 1):
        if T.joint_types[m] != U.JOINT_REVOLUTE or T.joint_types[m + 1] != U.JOINT_REVOLUTE:
            continue
        sysm = _reduced_system(T, c0, nr, m)
        if sysm is None:
            continue
        g, A, Bm, a1, a2, ucols, pr = sysm
        outs = [s for row in A for s in row] + [s for row in Bm for s in row] + a1 + a2
        live = g.live_lines(outs)
        cost = sum(ln.count("*") + ln.count(" + ") + ln.count(" - ") for ln in live)
        # what reduced_min_norm saves on structural zeros (reflector k that skips the last row: 11 - 2k FMAs; a zero of B: 2)
        z0 = lambda sc: sc[0] == "c" and sc[1] == 0.0
        sk = min(max(sum(z0(A[k][i]) for k in range(4)) for i in range(nu)), 3)
        eff = cost - sum(11 - 2 * k for k in range(sk)) - 2 * sum(z0(x) for row in Bm for x in row)
        if best is None or eff < best[0]:
            best = (eff, m, live, sysm, cost)
    if best is None:
        return stub
    _, m, live, (g, A, Bm, a1, a2, ucols, pr), cost = best
    # Sparsity the QR can use (the SKIP mechanism of min_norm_solve): put the unknown whose column of A has the most
    # structural zeros LAST and the task rows where it is zero FIRST; the first `skip` reflectors then leave it alone.
    is0 = lambda sc: sc[0] == "c" and sc[1] == 0.0
    zc = [sum(is0(A[k][i]) for k in range(4)) for i in range(nu)]
    last = max(range(nu), key=lambda i: (zc[i], i))
    skip = min(zc[last], 3)
    uperm = [i for i in range(nu) if i != last] + [last]
    kperm = [k for k in range(4) if is0(A[k][last])][:skip] + [k for k in range(4) if not is0(A[k][last])]
    kperm += [k for k in range(4) if k not in kperm]
    A = [[A[k][i] for i in uperm] for k in kperm]
    Bm = [[Bm[r][i] for i in uperm] for r in range(2)]
    a1 = [a1[k] for k in kperm]
    ucols = [ucols[i] for i in uperm]
    zb = sum(1 << (r * nu + i) for r in range(2) for i in range(nu) if is0(Bm[r][i]))
    out = [f"  // reduced path-A system in the frame of joint {m} (columns of joints {m}, {m + 1} eliminated; pivot row {pr}); ~{cost} fp32 ops",
           f"  // unknown order u = joints {ucols}; the last one is structurally zero in the first {skip} task rows",
           "  static constexpr bool kHasReduced = true;",
           f"  static constexpr int kRedC0 = {c0}, kRedNR = {nr}, kRedNU = {nu}, kRedM = {m - c0}, kRedSkip = {skip};",
           f"  static constexpr uint32_t kRedZB = 0x{zb:x}u;  // bit r * kRedNU + i: Bm[r][i] == 0 exactly",
           "  static __device__ __forceinline__ void reduced(const float (&q)[N], const float (&sn)[N], const float (&cs)[N], const float (&b)[6],",
           "      float (&At)[kRedNU][4], float (&Bm)[2][kRedNU], float (&a1)[4], float (&a2)[2]) {",
           "    (void)q; (void)sn; (void)cs;"]
    out += ["    " + ln for ln in live]
    for i in range(nu):
        out.append("    " + " ".join(f"At[{i}][{k}] = {g.expr(A[k][i])};" for k in range(4)))
    for r in range(2):
        out.append("    " + " ".join(f"Bm[{r}][{i}] = {g.expr(Bm[r][i])};" for i in range(nu)))
    out.append("    " + " ".join(f"a1[{k}] = {g.expr(a1[k])};" for k in range(4)))
    out.append("    " + " ".join(f"a2[{k}] = {g.expr(a2[k])};" for k in range(2)))
    out.append("  }")
    out += _gen_reduced_ik(T, c0, nr, m, uperm, kperm)
    # dq of the controlled range from (u, v): v[0] = joint m+1, v[1] = joint m
    out.append("  static __device__ __forceinline__ void reduced_scatter(const float (&u)[kRedNU], const float (&v)[2], float (&dq)[kRedNR]) {")
    asg = []
    for r in range(nr):
        j = c0 + r
        src = "v[1]" if j == m else ("v[0]" if j == m + 1 else f"u[{ucols.index(j)}]")
        asg.append(f"dq[{r}] = {src};")
    out.append("    " + " ".join(asg))
    out.append("  }")
    return out


def _gen_reduced_ik(T: U.ChainTable, c0: int, nr: int, m: int, uperm, kperm):
    """Path-B members for the same frame / permutations as ``reduced``: the iteration of the absolute-target solver
    (kinematics.py:144-155) needs the pose ERROR as its twist, and forming it in F_m shares the two half chains with
    the reduced system instead of running a second, base-frame FK:
      reduced_ik_sys(q, sn, cs, tp, Rt) -> At, Bm (as ``reduced``), dp = target position - EE position, Tm = target
          orientation, Cm = EE orientation, pe = EE position, all expressed in F_m;
      reduced_ik_rhs(dp, w, pe) -> a1, a2 for the twist (dp, w) at the EE (w = rotation error vector in F_m)."""
    nu = nr - 2
    fo = {}
    g, A, Bm, _, _, ucols, pr = _reduced_system(T, c0, nr, m, fo)
    c = g.c
    Rb, pe, Re, pb = fo["Rb"], fo["pe"], fo["Re"], fo["pb"]
    tp = [g.v(f"tp[{k}]") for k in range(3)]
    Rt = [[g.v(f"Rt[{3 * r + k}]") for k in range(3)] for r in range(3)]
    ptm = [g.sumprod([[Rb[r][k], tp[k]] for k in range(3)] + [[pb[r]]]) for r in range(3)]
    dp = [g.sumprod([[ptm[r]], [pe[r], c(-1.0)]]) for r in range(3)]
    Tm = [[g.sumprod([[Rb[r][k], Rt[k][j]] for k in range(3)]) for j in range(3)] for r in range(3)]
    A = [[A[k][i] for i in uperm] for k in kperm]
    Bm = [[Bm[r][i] for i in uperm] for r in range(2)]
    outs = [x for row in A for x in row] + [x for row in Bm for x in row] + dp + pe + [x for row in Tm for x in row] + [x for row in Re for x in row]
    live = g.live_lines(outs)
    cost = sum(ln.count("*") + ln.count(" + ") + ln.count(" - ") for ln in live)
    out = [f"  // path-B system in the same frame: reduced matrix + pose-error ingredients; ~{cost} fp32 ops",
           "  static __device__ __forceinline__ void reduced_ik_sys(const float (&q)[N], const float (&sn)[N], const float (&cs)[N],",
           "      const float (&tp)[3], const float (&Rt)[9], float (&At)[kRedNU][4], float (&Bm)[2][kRedNU], float (&dp)[3],",
           "      float (&Tm)[9], float (&Cm)[9], float (&pe)[3]) {",
           "    (void)q; (void)sn; (void)cs;"]
    out += ["    " + ln for ln in live]
    for i in range(nu):
        out.append("    " + " ".join(f"At[{i}][{k}] = {g.expr(A[k][i])};" for k in range(4)))
    for r in range(2):
        out.append("    " + " ".join(f"Bm[{r}][{i}] = {g.expr(Bm[r][i])};" for i in range(nu)))
    out.append("    " + " ".join(f"dp[{k}] = {g.expr(dp[k])};" for k in range(3)) + " " + " ".join(f"pe[{k}] = {g.expr(pe[k])};" for k in range(3)))
    for r in range(3):
        out.append("    " + " ".join(f"Tm[{3 * r + j}] = {g.expr(Tm[r][j])};" for j in range(3)) + " "
                   + " ".join(f"Cm[{3 * r + j}] = {g.expr(Re[r][j])};" for j in range(3)))
    out.append("  }")
    # right-hand side: twist at the EE -> reference point on joint m's axis -> the constant row operations
    g2 = Gen()
    fo2 = {}
    _reduced_system_rowop_only = _reduced_system(T, c0, nr, m, fo2)  # same constants; only its rowop closure is used
    g2 = _reduced_system_rowop_only[0]
    mark = len(g2.lines)
    dpv = [g2.v(f"dp[{k}]") for k in range(3)]
    wv = [g2.v(f"w[{k}]") for k in range(3)]
    pev = [g2.v(f"pe[{k}]") for k in range(3)]
    pxw = _cross(g2, pev, wv)
    v = [g2.sumprod([[dpv[r]], [pxw[r]]]) for r in range(3)]
    rr = fo2["rowop"](v + wv)
    K = [r for r in range(5) if r != pr]
    a1 = [rr[r] for r in K]
    a1 = [a1[k] for k in kperm]
    a2 = [rr[pr], rr[5]]
    live2 = [ln for ln in g2.live_lines(a1 + a2) if ln in g2.lines[mark:]]
    out += ["  static __device__ __forceinline__ void reduced_ik_rhs(const float (&dp)[3], const float (&w)[3], const float (&pe)[3],",
            "      float (&a1)[4], float (&a2)[2]) {"]
    out += ["    " + ln for ln in live2]
    out.append("    " + " ".join(f"a1[{k}] = {g2.expr(a1[k])};" for k in range(4)))
    out.append("    " + " ".join(f"a2[{k}] = {g2.expr(a2[k])};" for k in range(2)))
    out.append("  }")
    return out


# ----------------------------------------------------------------------------------------------------------
# 3-dof path A (pd_ee_delta_pos: J[:, 0:3], kinematics.py:169-170).  A rotation of the task frame is an invertible
# row operation too, so the linear Jacobian (reference point = EE origin, which the position task fixes) is
# generated in the joint frame F_m that makes it cheapest: the two half chains from joint m are shorter products
# than one chain from the base.  The dense NR x 3 Householder solve stays.
# ----------------------------------------------------------------------------------------------------------
def gen_linear(T: U.ChainTable, c0: int, nr: int):
    """Member lines of the chain struct: ``kHasLinear`` and ``linear(q, sn, cs, b, At[NR][3], a3[3])``."""
    n = T.n
    best = None
    first_rev = min((j for j in range(n) if T.joint_types[j] == U.JOINT_REVOLUTE), default=0)
    for m in range(max(c0 - 1, first_rev - 1, -1), c0 + nr):  # never before base prismatics: lever arms must stay small
        g = Gen()
        z, o, pe, Rb = _frame_chains(g, T, m)
        cols = []
        for j in range(c0, c0 + nr):
            if T.joint_types[j] == U.JOINT_REVOLUTE:
                d = [g.sumprod([[pe[r]], [o[j][r], g.c(-1.0)]]) for r in range(3)]
                cols.append(_cross(g, z[j], d))
            else:
                cols.append(list(z[j]))
        b = [g.v(f"b[{k}]") for k in range(3)]
        rhs = [g.sumprod([[Rb[r][k], b[k]] for k in range(3)]) for r in range(3)]
        outs = [x for col in cols for x in col] + rhs
        live = g.live_lines(outs)
        cost = sum(ln.count("*") + ln.count(" + ") + ln.count(" - ") for ln in live)
        if best is None or cost < best[0]:
            best = (cost, m, live, g, cols, rhs)
    cost, m, live, g, cols, rhs = best
    is0 = lambda sc: sc[0] == "c" and sc[1] == 0.0
    last_zero = all(is0(x) for x in cols[-1])
    out = [f"  // 3-dof (position) path-A system in the frame of joint {m} (-1 = base), reference point = EE; ~{cost} fp32 ops",
           "  static constexpr bool kHasLinear = true;",
           f"  static constexpr int kLinC0 = {c0}, kLinNR = {nr};",
           f"  static constexpr bool kLinLastZero = {'true' if last_zero else 'false'};  // the last controlled joint's column is exactly zero",
           "  static __device__ __forceinline__ void linear(const float (&q)[N], const float (&sn)[N], const float (&cs)[N], const float (&b)[3],",
           "      float (&At)[kLinNR][3], float (&a3)[3]) {",
           "    (void)q; (void)sn; (void)cs;"]
    out += ["    " + ln for ln in live]
    for r in range(nr):
        out.append("    " + " ".join(f"At[{r}][{k}] = {g.expr(cols[r][k])};" for k in range(3)))
    out.append("    " + " ".join(f"a3[{k}] = {g.expr(rhs[k])};" for k in range(3)))
    out.append("  }")
    return out
