"""Pin-on-arrival: the oracle (and, with a GPU, the kernels) against vectors dumped from the REAL pytorch_kinematics /
sapien by tools/gen_golden_pk.py.  Both packages are absent from this container and not installable (no network;
setup.py:46,50-51), so the fixture does not exist and these tests are skipped WITH THAT REASON; they become binding the
day `python tools/gen_golden_pk.py` finds the package (site-packages or baseline/_ref) and writes the fixture."""
import os

import numpy as np
import pytest

from conftest import FETCH, PANDA, ROOT
from oracle import oracle as O

FIX = os.path.join(ROOT, "tests", "golden", "pk_golden.npz")
pytestmark = pytest.mark.skipif(
    not os.path.exists(FIX),
    reason="pytorch_kinematics==0.7.5 / sapien absent: tests/golden/pk_golden.npz not generated (run tools/gen_golden_pk.py where they import)")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(FIX, allow_pickle=False))


@pytest.mark.parametrize("name,robot", [("panda", PANDA), ("fetch", FETCH)])
def test_oracle_jacobian_is_pk_chain_jacobian(gold, name, robot):
    """pk chain.jacobian(q, ret_eef_pose=True) (kinematics.py:157): rows [linear; angular], base frame, EE origin."""
    if f"{name}_jacobian" not in gold:
        pytest.skip("fixture has no pytorch_kinematics section")
    fr = O.load_chain(robot[0], robot[1])
    q = gold[f"{name}_q"].astype(np.float64)
    J, R, p = O.jacobian(fr, q)
    assert np.abs(J - gold[f"{name}_jacobian"]).max() < 5e-6  # pk computes in fp32
    M = gold[f"{name}_eef_matrix"]
    assert np.abs(M[:, :3, 3] - p).max() < 5e-6 and np.abs(M[:, :3, :3] - R).max() < 5e-6


def test_recalled_pseudo_inverse_ik_defaults(gold):
    """SURVEY A.2 values marked UNVERIFIED, against what the installed PseudoInverseIK really carries."""
    if "pik_solutions" not in gold:
        pytest.skip("fixture has no PseudoInverseIK section")
    from maniskill_b200.kinematics import Kinematics as K

    recalled = {"lr": K.lr, "pos_tolerance": K.pos_tolerance, "rot_tolerance": K.rot_tolerance, "max_iterations": K.max_iterations}
    for k, v in recalled.items():
        if f"pik_default_{k}" in gold:
            assert float(gold[f"pik_default_{k}"]) == pytest.approx(v), k
    for k in ("regularlization", "regularization"):
        if f"pik_default_{k}" in gold:
            assert float(gold[f"pik_default_{k}"]) == pytest.approx(K.regularization)


def test_oracle_iterative_ik_is_pk_solve(gold):
    """PseudoInverseIK.solve as called at kinematics.py:145-155 vs oracle.ik_iterative with the dumped defaults."""
    if "pik_solutions" not in gold:
        pytest.skip("fixture has no PseudoInverseIK section")
    fr = O.load_chain(PANDA[0], PANDA[1])
    tgt = np.concatenate([gold["pik_target_pos"], gold["pik_target_quat_wxyz"]], 1).astype(np.float64)
    kw = {}
    for k, name in (("lr", "lr"), ("pos_tolerance", "pos_tol"), ("rot_tolerance", "rot_tol")):
        if f"pik_default_{k}" in gold:
            kw[name] = float(gold[f"pik_default_{k}"])
    for k in ("regularlization", "regularization"):
        if f"pik_default_{k}" in gold:
            kw["reg"] = float(gold[f"pik_default_{k}"])
    ref = O.ik_iterative(fr, gold["pik_q0"].astype(np.float64), tgt, **kw)
    err = np.abs(ref - gold["pik_solutions"]).max(1)
    # fp32 library vs fp64 oracle: iteration flips allowed on a small fraction (see tests/test_gpu_kinematics.py)
    assert np.median(err) < 5e-6 and (err > 1e-4).mean() < 0.05, (np.median(err), err.max())
