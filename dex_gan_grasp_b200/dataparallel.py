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


def allreduce_gradients(module):
    """Average the gradients of `module` over all ranks with a single flat allreduce."""
    if not is_parallel():
        return
    params = [p for p in module.parameters() if p.grad is not None]
    if not params:
        return
    flat = torch.cat([p.grad.reshape(-1) for p in params])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
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
