"""Data-parallel host logic: the reference sums per-sample gradients over the batch
(src/SoftmaxCrossEntropyNeuralNet.java:305-322), so G ranks each take a contiguous slice of every
global mini-batch, run forward/backward locally, and one all-reduce(sum) per weight matrix gives the
single-device gradient; the update divides by the GLOBAL batch size (S:333, `batch.size()`).

One process per GPU (torchrun); `torch.distributed` is plumbing only (rendezvous + shipping the
NCCL unique id); the gradient all-reduce itself is issued by libb2nn.so on its own stream.
"""
import os

import numpy as np


def shard_bounds(global_B, rank, world):
    """rows [lo, hi) of a global mini-batch owned by `rank` (contiguous, sizes differ by at most 1)"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world %d/%d" % (rank, world))
    if global_B < world:
        raise ValueError("global batch %d smaller than world size %d" % (global_B, world))
    base, rem = divmod(global_B, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_index_lists(idx, rank, world):
    """idx [iters][global_B] (the sampler's draws, identical on every rank) -> this rank's
    [iters][B_local] slice.  Requires world | global_B so every rank launches the same shapes."""
    idx = np.asarray(idx)
    if idx.ndim != 2:
        raise ValueError("idx must be [iters][global_B]")
    if idx.shape[1] % world != 0:
        raise ValueError("global batch %d is not divisible by world size %d" % (idx.shape[1], world))
    lo, hi = shard_bounds(idx.shape[1], rank, world)
    return np.ascontiguousarray(idx[:, lo:hi])


def init_comm(net, dist, rank=None, world=None):
    """create the NCCL communicator inside libb2nn.so: rank 0 makes the unique id, the launcher's
    process group (any backend) ships it."""
    from .nets import commUniqueId
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    uid = [commUniqueId() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    net.commInit(uid[0], rank, world)
    if world > 1 and os.environ.get("B2NN_DP", "p2p") != "nccl":
        # peer-memory path: all-gather the CUDA IPC handles of the parameter arenas (plumbing only)
        handles = [None] * world
        dist.all_gather_object(handles, net.commIpcExport())
        net.commIpcImport(handles)
        dist.barrier()
    return rank, world
