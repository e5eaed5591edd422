"""oracle/part_filter_oracle.py against the reference's own functions (executed from /root/reference when present) and
known answers."""
import ast
import os

import numpy as np
import pytest

from oracle import part_filter_oracle as po

REF = "/root/reference/DexGanGrasp/utils/filter_grasps_given_mask.py"


def _case(seed=0, n=200, m=3000):
    rng = np.random.default_rng(seed)
    grasps = {"transl": 0.15 * rng.standard_normal((n, 3)), "joint_conf": rng.uniform(0, 1, (n, 12)),
              "rot_matrix": rng.standard_normal((n, 3, 3))}
    part = 0.1 * rng.standard_normal((m, 3)) + np.array([0.4, -0.2, 0.7])
    part[::50] += 3.0  # outliers removed by the 0.2 crop
    return grasps, part, np.array([0.38, -0.21, 0.72])


def test_known_answer():
    grasps = {"transl": np.array([[0.0, 0, 0], [1.0, 0, 0], [0.1, 0, 0]])}
    part = np.array([[1.1, 2.0, 3.0], [1.1, 2.0, 3.0], [9.0, 9.0, 9.0]])
    idx, mean, mins = po.filter_grasps_given_part(grasps, part, np.array([1.0, 2.0, 3.0]))
    assert np.allclose(mean, [[0.1, 0, 0]]) and list(idx) == [2, 0, 1] and np.allclose(mins, [0.01, 0.81, 0.0])


def test_against_reference_functions():
    if not os.path.exists(REF):
        pytest.skip("reference tree not on this machine")
    ns = {"np": np}
    for node in ast.parse(open(REF).read()).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("euclidean_distance_points_pairwise_np", "sort_grasps"):
            exec(compile(ast.Module([node], []), REF, "exec"), ns)
    grasps, part, center = _case()
    idx, mean, mins = po.filter_grasps_given_part(grasps, part, center)
    ref_dist = ns["euclidean_distance_points_pairwise_np"](grasps["transl"], mean)
    assert np.allclose(po.pairwise_sq(grasps["transl"], mean), ref_dist, rtol=1e-13)
    assert np.array_equal(np.argsort(np.min(ref_dist, axis=1)), idx)
    a = ns["sort_grasps"]({k: v.copy() for k, v in grasps.items()}, idx, 30)
    b = po.sort_grasps(grasps, idx, 30)
    assert all(np.array_equal(a[k], b[k]) for k in a)
