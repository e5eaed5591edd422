"""One-process-per-GPU data parallelism for the training step (replaces the reference's non-functional
torch.nn.DataParallel, DexGanGrasp/models/dexgangrasp.py:127-129).

Weights are replicated, the batch is sharded by rank, and after each backward ONE allreduce runs over a flat
contiguous gradient buffer per network (NCCL over NVLink on the B200 box, gloo in the CPU tests).  Train-mode
BatchNorm statistics are reduced across ranks inside train_ops.batch_norm so that the N-rank step equals the 1-rank
step on the concatenated batch.
"""
import torch
import torch.distributed as dist


def is_parallel():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


# Every collective of the training step goes through all_reduce_sum, which counts calls and bytes per kind ("grad",
# "bn_fwd", "bn_bwd"): bench.py reports them per step (the host counters also advance during a CUDA-graph capture, which is
# how CapturedTrainStep knows what one replay holds).
# "p2p_calls": exchanges done INSIDE a kernel over NVLink peer memory (train_engine.py), no NCCL launch
_COUNTERS = {"grad_calls": 0, "grad_bytes": 0, "bn_fwd_calls": 0, "bn_fwd_bytes": 0, "bn_bwd_calls": 0, "bn_bwd_bytes": 0, "p2p_calls": 0}


def collective_counters():
    return dict(_COUNTERS)


def all_reduce_sum(t, kind="grad", async_op=False):
    """Sum `t` over all ranks in place (no-op on one rank)."""
    if not is_parallel():
        return None
    _COUNTERS[kind + "_calls"] += 1
    _COUNTERS[kind + "_bytes"] += t.numel() * t.element_size()
    return dist.all_reduce(t, op=dist.ReduceOp.SUM, async_op=async_op)


def all_gather(out, t, kind="bn_fwd"):
    """out [world, ...] <- every rank's t"""
    _COUNTERS[kind + "_calls"] += 1
    _COUNTERS[kind + "_bytes"] += out.numel() * out.element_size()
    return dist.all_gather_into_tensor(out, t)


def allreduce_gradients(module):
    """Average the gradients of `module` over all ranks with a single flat allreduce."""
    if not is_parallel():
        return
    flat = getattr(module, "_dexgg_flat_grad", None)
    if flat is not None:  # train_engine: every .grad is a view of one flat buffer -> exchange it in place (no cat, no copy-back)
        all_reduce_sum(flat, "grad")
        flat.div_(dist.get_world_size())
        return
    params = [p for p in module.parameters() if p.grad is not None]
    if not params:
        return
    flat = torch.cat([p.grad.reshape(-1) for p in params])
    all_reduce_sum(flat, "grad")
    flat.div_(dist.get_world_size())
    off = 0
    for p in params:
        n = p.grad.numel()
        p.grad.copy_(flat[off:off + n].view_as(p.grad))
        off += n


def shard(batch, rank=None, world=None):
    """Rows [rank*B/world, (rank+1)*B/world) of every tensor of a batch dict."""
    if rank is None:
        rank, world = (dist.get_rank(), dist.get_world_size()) if is_parallel() else (0, 1)
    out = {}
    for k, v in batch.items():
        if torch.is_tensor(v):
            n = v.shape[0] // world
            out[k] = v[rank * n:(rank + 1) * n]
        else:
            out[k] = v
    return out


def shard_objects(n_objects, rank=None, world=None):
    """Inference sharding: rank r owns objects [lo, hi) with ALL their latent samples, so the per-object
    BPS projection stays rank-local and no collective is needed.  Remainders go to the first ranks."""
    if rank is None:
        rank, world = (dist.get_rank(), dist.get_world_size()) if is_parallel() else (0, 1)
    base, rem = divmod(int(n_objects), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def object_seed(seed, object_id):
    """Seed of object `object_id`'s latent stream: results do not depend on how objects are spread over GPUs."""
    return (int(seed) * 1000003 + int(object_id)) & 0x7FFFFFFF


def batchnorm_statistics(x, eps=1e-5):
    """The two-pass cross-rank statistics protocol of train_ops._BatchNormFn, on plain tensors (any device): allreduce
    the column sums -> global mean, allreduce the centred sums of squares -> biased variance.  Returns (mean, invstd,
    unbiased variance, global row count).  train_ops runs the same sequence with the CUDA reduction kernels."""
    world = dist.get_world_size() if is_parallel() else 1
    count = float(x.shape[0] * world)
    s = x.sum(0)
    if world > 1:
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    mean = s / count
    ss = ((x - mean) ** 2).sum(0)
    if world > 1:
        dist.all_reduce(ss, op=dist.ReduceOp.SUM)
    var = ss / count
    return mean, 1.0 / torch.sqrt(var + eps), var * (count / max(count - 1.0, 1.0)), count
