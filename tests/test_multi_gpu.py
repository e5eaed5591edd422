"""Data-parallel training step on two real GPUs (skipped on a one-GPU box): tools/train_dp_check.py under torchrun - every rank first
runs the full-batch step alone, then the ranks run the sharded step (BatchNorm statistics and gradients exchanged inside our kernels
over NVLink peer mappings, no NCCL call in the step); losses, updated weights and BatchNorm statistics must agree with the one-process
step on the concatenated batch and the ranks must end bit-identical.  Covers the bf16 fused engine (eager and captured, GAN +
evaluator as two concurrent chains) and the WGAN-GP variant's twice-differentiable cross-rank BatchNorm."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    env = dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS="32", **env_extra)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(REPO, "tools", "train_dp_check.py")]
    out = subprocess.run(cmd, env=env, cwd=REPO, capture_output=True, text=True, timeout=600)
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert out.returncode == 0 and lines, out.stdout[-2000:] + out.stderr[-2000:]
    return json.loads(lines[-1])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("graph", ["0", "1"])
def test_two_rank_engine_step_equals_one_rank_step(graph):
    r = _run({"DP_BATCH": "1024", "DEXGG_PRECISION": "bf16", "DP_COMBINED": "1", "DP_GRAPH": graph})
    assert r["world"] == 2 and "fused" in r["engine"]
    assert r["collectives"]["grad_calls"] == 0 and r["collectives"]["p2p_calls"] > 0  # nothing through NCCL
    assert r["max_rank_divergence"] == 0.0
    assert max(r["loss_abs_diff"].values()) < 5e-4, r["loss_abs_diff"]
    assert r["max_weight_diff"] <= 2.1e-4 * (2 if graph == "1" else 1) and r["max_buffer_diff"] < 2e-3, r
    assert r["worst_grad_rel_diff"][0] < 0.3, r["worst_grad_rel_diff"]  # bf16: LeakyReLU slope flips (tolerances.py)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_wgan_gp_step_equals_one_rank_step():
    r = _run({"DP_BATCH": "512", "DEXGG_PRECISION": "fp32", "DP_WGAN": "1"})
    assert r["wgan"] and r["max_rank_divergence"] == 0.0
    assert max(r["loss_abs_diff"].values()) < 1e-5, r["loss_abs_diff"]
    assert r["worst_grad_rel_diff"][0] < 5e-3 and r["max_weight_diff"] <= 2.1e-4, r
