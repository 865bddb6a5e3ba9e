"""Data-parallel plumbing (one process per GPU, torch.distributed): what the reference's DDP wrapper was meant to
do (train.py:181-196,230-235) but never did (its forward bypasses `ddp_model`).

All parameters (and all gradients) of an XUNet are views of ONE flat fp32 arena, so the collective work of a step is
a single all-reduce of one tensor; NCCL's NVLS/ring does the rest over NVLink.  No data-path collective exists in
sampling (objects are sharded, `sampling.shard_objects`)."""
import torch
import torch.distributed as dist


def world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def rank():
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def broadcast_parameters(flat, src=0):
    """same initial weights on every rank (DDP's constructor broadcast)"""
    if world() > 1:
        dist.broadcast(flat, src)


def allreduce_mean_(flat, async_op=False):
    """in-place mean over ranks of the flat gradient arena (each rank's loss is the mean over its own shard, so the
    mean of the per-rank gradients is the gradient of the global-batch mean loss)"""
    if world() == 1:
        return None
    if dist.get_backend() == "nccl":
        return dist.all_reduce(flat, op=dist.ReduceOp.AVG, async_op=async_op)
    w = dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=async_op)   # gloo has no AVG
    if async_op:
        w.wait()
    flat.div_(world())
    return None


def allreduce_sum_(flat):
    """in-place SUM over ranks (the caller folds the 1/world into its consumer, e.g. Adam's grad_scale: a plain sum keeps every
    NCCL algorithm eligible, an AVG is a pre-multiplied sum)"""
    if world() > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)


def shard_batch(batch, rank_, world_):
    """contiguous shard of a global batch (tuple of tensors with leading batch dim) for this rank"""
    B = batch[0].shape[0]
    assert B % world_ == 0, f"global batch {B} not divisible by {world_} ranks"
    per = B // world_
    return tuple(t[rank_ * per:(rank_ + 1) * per] for t in batch)
