"""Data-parallel host logic: the reference sums per-sample gradients over the batch
(src/SoftmaxCrossEntropyNeuralNet.java:305-322), so G ranks each take a contiguous slice of every
global mini-batch, run forward/backward locally, and one all-reduce(sum) per weight matrix gives the
single-device gradient; the update divides by the GLOBAL batch size (S:333, `batch.size()`).

One process per GPU (torchrun); `torch.distributed` is plumbing only (rendezvous + shipping the NCCL unique id, the
CUDA IPC handles or the (pid, fd) pairs of the multicast set-up); the exchange itself -- by default ONE fused
reduce-scatter + momentum update + all-gather kernel per layer over NVLink peer memory, summing inside the NVSwitch from
4 GPUs up; NCCL all-reduce + update as the baseline -- is issued by libb2nn.so on its own stream.
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
    # B2NN_DP = nvls | p2p | nccl.  Default: the in-switch reduction from 4 GPUs up (measured at 8 GPUs: step
    # 0.769 ms against 0.848 ms for plain peer memory), plain peer memory at 2 (there multimem sends the local
    # half through the switch as well: 0.799 against 0.768 ms).  profiles/r01_scaling.md has the runs.
    mode = os.environ.get("B2NN_DP") or ("nvls" if world >= 4 else "p2p")
    # (TF32 nets take the same exchanges: the fused kernel replicates their FP32 weights instead of a BF16 copy)
    if world > 1 and mode == "nvls":
        if _init_nvls(net, dist, rank, world):
            return rank, world
        mode = "p2p"  # no multicast on this box: nothing was changed, use plain peer memory
    if world > 1 and mode != "nccl":
        # peer-memory path: all-gather the CUDA IPC handles of the parameter arenas (plumbing only)
        handles = [None] * world
        dist.all_gather_object(handles, net.commIpcExport())
        net.commIpcImport(handles)
        dist.barrier()
    return rank, world


def _agree(dist, world, ok, what):
    """every rank learns whether `what` succeeded everywhere, so that all ranks take the same branch (a rank that
    raised alone would leave the others waiting in the next collective)"""
    flags = [None] * world
    dist.all_gather_object(flags, None if ok is True else str(ok))
    bad = [(r, f) for r, f in enumerate(flags) if f is not None]
    if bad:
        raise RuntimeError("NVSwitch multicast set-up failed in %s on rank(s) %s: %s (set B2NN_DP=p2p to use plain "
                           "peer memory)" % (what, [r for r, _ in bad], bad[0][1]))


def _init_nvls(net, dist, rank, world):
    """NVSwitch multicast set-up (include/b2nn.h, b2nn_comm_nvls_*): only (pid, fd) pairs travel through
    the process group.  False (on every rank) when any rank lacks multicast support; RuntimeError on EVERY rank
    when a later stage fails anywhere (the arena has moved by then: re-create the net to fall back)."""
    mine = net.commNvlsExport()
    pairs = [None] * world
    dist.all_gather_object(pairs, mine)
    if all(p is None for p in pairs):
        return False
    if any(p is None for p in pairs):
        # every rank sees the same gathered list, so every rank raises (a rank raising alone would leave the others
        # waiting in the next collective); the ranks that did export have moved their arena: re-create the net
        raise RuntimeError("NVSwitch multicast support differs between ranks: rank(s) %s lack it (set B2NN_DP=p2p)"
                           % [r for r, p in enumerate(pairs) if p is None])

    def stage(what, fn):
        try:
            out, ok = fn(), True
        except Exception as e:  # noqa: BLE001 -- reported on all ranks by _agree
            out, ok = None, "%s: %s" % (type(e).__name__, e)
        _agree(dist, world, ok, what)
        return out

    stage("import of the peers' arenas", lambda: net.commNvlsImport(pairs))
    team = [stage("creation of the multicast object", lambda: net.commNvlsTeam(None) if rank == 0 else None)]
    dist.broadcast_object_list(team, src=0)   # rank 0 keeps its fd open until everybody has duplicated it
    stage("joining the multicast team", lambda: net.commNvlsTeam(team[0]) if rank != 0 else None)
    stage("binding the arena", lambda: net.commNvlsBind())   # _agree doubles as the barrier before / after the bind
    return True
