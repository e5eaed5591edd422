"""Host-side logic of the N>1 paths on CPU with the gloo backend, world_size 2:
  * inference: objects shard over ranks with no collective; per-object latent seeds make results rank-count independent;
  * training: flat gradient allreduce (average) and the two-pass BatchNorm statistics protocol give the 1-rank result on
    the concatenated batch.
The CUDA kernels themselves cannot run here; what is tested is the exchange protocol around them."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dex_gan_grasp_b200 import dataparallel as dp

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, port, fn, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        out[rank] = fn(rank)
    finally:
        dist.destroy_process_group()


def _run(fn):
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(_free_port(), fn, out), nprocs=WORLD, join=True)
    return [out[r] for r in range(WORLD)]


def _full_batch():
    g = torch.Generator().manual_seed(0)
    return torch.randn(64, 37, generator=g) * 3 + 1.5


def _grad_job(rank):
    torch.manual_seed(0)
    lin = torch.nn.Linear(37, 5)
    x = dp.shard({"x": _full_batch()}, rank, WORLD)["x"]
    lin(x).pow(2).mean().backward()
    dp.allreduce_gradients(lin)
    return [p.grad.clone() for p in lin.parameters()]


def test_gradient_allreduce_equals_full_batch():
    res = _run(_grad_job)
    torch.manual_seed(0)
    lin = torch.nn.Linear(37, 5)
    lin(_full_batch()).pow(2).mean().backward()
    for r in range(WORLD):
        for got, p in zip(res[r], lin.parameters()):
            assert torch.allclose(got, p.grad, atol=1e-6, rtol=1e-5)


def _bn_job(rank):
    x = dp.shard({"x": _full_batch()}, rank, WORLD)["x"]
    mean, invstd, unbiased, count = dp.batchnorm_statistics(x)
    return mean, invstd, unbiased, count


def test_batchnorm_statistics_equal_full_batch():
    res = _run(_bn_job)
    x = _full_batch()
    bn = torch.nn.BatchNorm1d(37, momentum=1.0)  # momentum 1: running stats = this batch's statistics
    bn.train()
    y = bn(x)
    for mean, invstd, unbiased, count in res:
        assert count == 64
        assert torch.allclose(mean, bn.running_mean, atol=1e-6)
        assert torch.allclose(unbiased, bn.running_var, rtol=1e-5, atol=1e-6)
        assert torch.allclose((x - mean) * invstd, y, atol=1e-5)


def _penalty_grads(x, lin, bn):
    """A WGAN-GP-shaped second-order quantity through the twice-differentiable BatchNorm: the input gradient of the summed scores,
    its per-row norm penalty, differentiated w.r.t. the parameters."""
    from dex_gan_grasp_b200 import train_ops
    x = x.clone().requires_grad_(True)
    s = torch.sigmoid(train_ops._batch_norm_dd(lin(x), bn).sum(1, keepdim=True))
    gx = torch.autograd.grad(s, x, torch.ones_like(s), create_graph=True)[0]
    penalty = ((gx.norm(2, dim=1) - 1) ** 2).mean()
    params = [lin.weight, lin.bias, bn.weight, bn.bias]
    return penalty.detach(), torch.autograd.grad(penalty, params)


def _make_lin_bn():
    torch.manual_seed(3)
    lin = torch.nn.Linear(37, 11)
    bn = torch.nn.BatchNorm1d(11)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    bn.train()
    return lin, bn


def _bn_dd_job(rank):
    lin, bn = _make_lin_bn()
    x = dp.shard({"x": _full_batch()}, rank, WORLD)["x"]
    pen, grads = _penalty_grads(x, lin, bn)
    grads = [g.clone() for g in grads]
    for g in grads:  # what allreduce_gradients does with .grad: average over the ranks
        dist.all_reduce(g)
        g.div_(WORLD)
    dist.all_reduce(pen)
    return pen / WORLD, grads, bn.running_mean.clone(), bn.running_var.clone()


def test_double_backward_batchnorm_two_ranks_equals_full_batch():
    """train_ops._batch_norm_dd with cross-rank statistics (_AllReduceSumFn, whose backward is the same all-reduce): the penalty and
    its parameter gradients on two ranks equal the one-process result on the concatenated batch, as do the running statistics."""
    res = _run(_bn_dd_job)
    lin, bn = _make_lin_bn()
    pen, grads = _penalty_grads(_full_batch(), lin, bn)
    for r in range(WORLD):
        p_r, g_r, rm, rv = res[r]
        assert torch.allclose(p_r, pen, rtol=1e-5, atol=1e-7)
        for a_, b_ in zip(g_r, grads):
            assert torch.allclose(a_, b_, rtol=1e-4, atol=1e-6), float((a_ - b_).abs().max())
        assert torch.allclose(rm, bn.running_mean, atol=1e-6) and torch.allclose(rv, bn.running_var, rtol=1e-5, atol=1e-6)


def _shard_job(rank):
    lo, hi = dp.shard_objects(1024 + 3, rank, WORLD)
    seeds = [dp.object_seed(7, o) for o in range(lo, hi)]
    return lo, hi, seeds


def test_object_sharding_covers_all_objects_once():
    res = _run(_shard_job)
    (lo0, hi0, s0), (lo1, hi1, s1) = res
    assert lo0 == 0 and hi0 == lo1 and hi1 == 1027 and (hi0 - lo0) - (hi1 - lo1) in (0, 1)
    assert s0 + s1 == [dp.object_seed(7, o) for o in range(1027)]  # same per-object latent streams as a 1-rank run
    assert len(set(s0 + s1)) == 1027


def test_single_process_is_a_noop():
    assert not dp.is_parallel()
    assert dp.shard_objects(10) == (0, 10)
    lin = torch.nn.Linear(3, 2)
    lin(torch.ones(4, 3)).sum().backward()
    g = lin.weight.grad.clone()
    dp.allreduce_gradients(lin)
    assert torch.equal(g, lin.weight.grad)
    with pytest.raises(Exception):
        dist.get_rank()
