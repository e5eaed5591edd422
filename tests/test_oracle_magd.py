"""The MAGD restatement (oracle/magd_oracle.py) against golden vectors produced by executing the reference's own
functions (eval.py:48-137; tests/golden/make_magd_golden.py)."""
import os

import numpy as np
import pytest

from oracle import magd_oracle


def cases(g):
    for i in range(int(g["n_cases"])):
        pred = {k: g[f"c{i}_pred_{k}"] for k in ("transl", "rot_matrix", "joint_conf")}
        gt = {k: g[f"c{i}_gt_{k}"] for k in ("transl", "rot_matrix", "joint_conf")}
        yield i, pred, gt, g[f"c{i}_result"]


@pytest.fixture(scope="module")
def magd_golden():
    from conftest import GOLDEN
    return np.load(os.path.join(GOLDEN, "magd_golden.npz"))


def test_oracle_matches_reference_golden(magd_golden):
    n = 0
    for i, pred, gt, want in cases(magd_golden):
        got = magd_oracle.magd(pred, gt)
        assert np.allclose(got, want, rtol=1e-12, atol=1e-12), (i, got, want)
        n += 1
    assert n == 6


def test_identical_sets_have_zero_deviation_and_full_coverage(magd_golden):
    _, _, gt, _ = next(cases(magd_golden))
    t, r, j, cov = magd_oracle.magd(gt, gt)
    assert t == 0.0 and j == 0.0 and r < 1e-6 and cov == 1.0


def test_live_reference_when_present(magd_golden):
    if not os.path.exists("/root/reference/eval.py"):
        pytest.skip("reference tree not on this machine")
    import ast
    import torch
    ns = {"np": np, "torch": torch}
    for node in ast.parse(open("/root/reference/eval.py").read()).body:
        if isinstance(node, ast.FunctionDef) and ("pairwise_np" in node.name or node.name == "magd_for_grasp_distribution"):
            exec(compile(ast.Module([node], []), "eval.py", "exec"), ns)
    for i, pred, gt, want in cases(magd_golden):
        res = ns["magd_for_grasp_distribution"]({k: v.copy() for k, v in pred.items()}, {k: v.copy() for k, v in gt.items()})
        assert np.allclose(res, want, rtol=1e-12), i
