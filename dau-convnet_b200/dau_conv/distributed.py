"""Batch-sharded data parallelism for the DAU layer: only the parameter gradients cross GPUs.

The reference has no multi-GPU support; this is the (e) row of the scope
table: rank r owns images [r*N/R, (r+1)*N/R), parameters are replicated, and the four
[1,S,G,F] gradient tensors are summed with ONE all-reduce of a flat [4,S,G,F] buffer
(NCCL over NVLink on GPUs; gloo in the CPU tests of the host logic).
"""
import torch
import torch.distributed as dist


def shard_range(n_images, rank, world):
    """Contiguous image range of `rank` (the first n_images % world ranks get one more)."""
    base, extra = divmod(n_images, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class GradBucket(object):
    """One flat buffer holding dw, dmu1, dmu2, dsigma as views; all-reduced in one call."""

    def __init__(self, param_shape, device, dtype=torch.float32):
        self.flat = torch.zeros((4,) + tuple(param_shape), device=device, dtype=dtype)

    @property
    def views(self):
        return [self.flat[k] for k in range(4)]

    def zero_(self):
        self.flat.zero_()

    def all_reduce(self, group=None, async_op=False):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        return None


def all_reduce_param_grads(grads, group=None):
    """Sum a list of same-shaped gradient tensors across ranks with one collective."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return grads
    flat = torch.stack([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    for g, f in zip(grads, flat):
        g.copy_(f.view_as(g))
    return grads
