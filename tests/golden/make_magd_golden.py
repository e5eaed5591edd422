#!/usr/bin/env python
"""Generates tests/golden/magd_golden.npz by EXECUTING the reference's own MAGD functions (eval.py:48-137).

eval.py cannot be imported as a module here (open3d, tensorboard, dataset classes at import time), so the four function
definitions are extracted from /root/reference/eval.py with `ast` and executed unmodified in a namespace that holds only
numpy and torch.  No reference source is copied into the repository.  Run in the build container only.
Cases: (a) the 64 ground-truth grasps of data/eval_batch.pth against themselves perturbed (seeded), grouped per object;
(b) random sets of different sizes, f32 predictions against f64 ground truth as eval.py:309-369 feeds them."""
import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
from oracle import magd_oracle  # noqa: E402

src = open("/root/reference/eval.py").read()
want = {"geodesic_distance_rotmats_pairwise_np", "euclidean_distance_points_pairwise_np",
        "euclidean_distance_joint_conf_pairwise_np", "magd_for_grasp_distribution"}
tree = ast.parse(src)
ns = {"np": np, "torch": torch}
for node in tree.body:
    if isinstance(node, ast.FunctionDef) and node.name in want:
        exec(compile(ast.Module([node], []), "/root/reference/eval.py", "exec"), ns)
assert want <= set(ns)
ref_magd = ns["magd_for_grasp_distribution"]


def rand_rot(n, rng):
    q, r = np.linalg.qr(rng.standard_normal((n, 3, 3)))
    q = q * np.sign(np.diagonal(r, axis1=1, axis2=2))[:, None, :]
    q[:, :, 0] *= np.sign(np.linalg.det(q))[:, None]
    return q


out = {}
rng = np.random.default_rng(5)
eb = np.load(os.path.join(HERE, "eval_batch.npz"), allow_pickle=True)
cases = []
gt = {"transl": eb["transl"], "rot_matrix": eb["rot_matrix"], "joint_conf": eb["joint_conf"]}
pred = {"transl": (gt["transl"] + 0.01 * rng.standard_normal(gt["transl"].shape)).astype(np.float32),
        "rot_matrix": (gt["rot_matrix"] @ rand_rot(64, rng) if False else np.einsum("nij,njk->nik", gt["rot_matrix"], rand_rot(64, rng))).astype(np.float32),
        "joint_conf": (gt["joint_conf"] + 0.05 * rng.standard_normal(gt["joint_conf"].shape)).astype(np.float32)}
cases.append((pred, gt))
for n, m in ((1, 1), (100, 17), (1000, 64), (333, 500)):
    cases.append(({"transl": (0.1 * rng.standard_normal((n, 3))).astype(np.float32), "rot_matrix": rand_rot(n, rng).astype(np.float32),
                   "joint_conf": rng.uniform(0, 0.5, (n, 12)).astype(np.float32)},
                  {"transl": 0.1 * rng.standard_normal((m, 3)), "rot_matrix": rand_rot(m, rng), "joint_conf": rng.uniform(0, 0.5, (m, 12))}))
# duplicated ground-truth rows: np.argmin must take the first one
p, g = cases[2]
g2 = {k: np.concatenate([v, v[:5]]) for k, v in g.items()}
cases.append((p, g2))
for i, (p, g) in enumerate(cases):
    res = ref_magd({k: v.copy() for k, v in p.items()}, {k: v.copy() for k, v in g.items()})
    ours = magd_oracle.magd(p, g)
    assert np.allclose(res, ours, rtol=1e-12, atol=1e-12), (i, res, ours)
    for k in p:
        out[f"c{i}_pred_{k}"] = p[k]
        out[f"c{i}_gt_{k}"] = g[k]
    out[f"c{i}_result"] = np.asarray(res, np.float64)
out["n_cases"] = np.asarray(len(cases))
np.savez_compressed(os.path.join(HERE, "magd_golden.npz"), **out)
print("wrote magd_golden.npz;", len(cases), "cases; oracle == reference")
