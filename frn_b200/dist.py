"""Batch sharding of the fusion block across GPUs.

Every op of the block is per-sample (instance statistics over H*W, attention
inside one sample), so the only exchange is the sum of the fusion parameters'
gradients: one allreduce per step over a flat bucket (16 729 920 fp32 = 66.9 MB
for the four DSEC levels).  The reference's equivalent is nn.DataParallel's
reduce-to-device-0 (train_dsec.py:113).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_batch(global_batch: int, rank: int, world: int):
    """Contiguous slice [lo, hi) of the batch owned by `rank` (remainder to the low ranks)."""
    base, rem = divmod(global_batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class GradBucket:
    """Flat fp32 bucket aliasing the .grad of every parameter, so the allreduce is one call
    and needs no pack/unpack copies.

    Build it once and reuse it.  Between steps clear the gradients with ``bucket.zero()`` or
    ``optimizer.zero_grad(set_to_none=False)``: both keep the aliasing.  The reference loop's
    ``optimizer.zero_grad()`` (train_dsec_det_fast.py:237) defaults to ``set_to_none=True`` in torch >= 2.0,
    which detaches every ``.grad`` from the bucket; ``allreduce()`` therefore re-attaches first (``attach()``):
    a gradient that no longer lives in the bucket is copied in and ``p.grad`` is pointed at its view again, a
    parameter without a gradient contributes zeros.  That costs one copy per detached parameter but can never
    reduce a stale buffer."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.views, off = [], 0
        for p in self.params:
            n = p.numel()
            self.views.append(self.flat[off:off + n].view_as(p))
            off += n
        self.attach()

    def attach(self):
        """Make every p.grad the bucket's view (copying gradients produced outside the bucket).  Returns the
        number of parameters that had to be re-attached."""
        moved = 0
        for p, view in zip(self.params, self.views):
            g = p.grad
            if g is not None and g.data_ptr() == view.data_ptr() and g.dtype == view.dtype:
                continue
            if g is None:
                view.zero_()
            else:
                view.copy_(g)
            p.grad = view
            moved += 1
        return moved

    def zero(self):
        self.flat.zero_()
        self.attach()

    def allreduce(self, average: bool = True, async_op: bool = False):
        self.attach()
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return None
        if average:
            self.flat.div_(dist.get_world_size())
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=async_op)


class LevelBuckets:
    """One GradBucket per REFusion module (FPN level): 0.79 / 3.2 / 12.6 / 50.4 MB at the four DSEC levels
    (SURVEY.md 8.2 e).  ``allreduce_level(k)`` is called right after level k's backward has been enqueued -- on the
    stream that ran it -- with ``async_op=True``: NCCL orders the collective after that stream's work and runs it on its
    own stream, so the reduction of one level's gradients overlaps the backward of the levels still running.
    ``wait()`` makes the current stream wait for all of them.  Every rank must call the levels in the same order."""

    def __init__(self, modules):
        self.buckets = [GradBucket(list(m.parameters())) for m in modules]
        self._pending = []

    @property
    def numel(self):
        return sum(b.flat.numel() for b in self.buckets)

    def zero(self):
        for b in self.buckets:
            b.zero()

    def allreduce_level(self, k: int, average: bool = True):
        work = self.buckets[k].allreduce(average=average, async_op=True)
        if work is not None:
            self._pending.append(work)
        return work

    def wait(self):
        for w in self._pending:
            w.wait()
        self._pending = []

    def grad_norm(self):
        return torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(b.flat) for b in self.buckets]))


def allreduce_fusion_grads(params, average: bool = True):
    """One-shot allreduce of the given parameters' gradients (packs, reduces, unpacks).
    For the zero-copy path build a GradBucket once and reuse it."""
    params = [p for p in params if p.grad is not None]
    if not params or not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return
    flat = torch.cat([p.grad.reshape(-1).to(torch.float32) for p in params])
    if average:
        flat.div_(dist.get_world_size())
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    off = 0
    for p in params:
        n = p.numel()
        p.grad.copy_(flat[off:off + n].view_as(p.grad))
        off += n
