"""Worker bodies for the multi-process tests (spawned by test_dp.py).  The CPU (gloo) worker checks
the data-parallel HOST logic against the oracle; the GPU (nccl) worker checks the CUDA path."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _init(rank, world, port, backend):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group(backend, rank=rank, world_size=world)
    return dist


def gloo_worker(rank, world, port, out_dir):
    """Every rank draws the same global index lists, takes its slice (dp.shard_index_lists), sums the
    oracle's per-sample gradients over the slice, all-reduces the sums over gloo and applies the
    reference's update with the GLOBAL batch size; rank 0 stores the result for the parent to compare
    with a single-process oracle run."""
    import torch
    import __graft_entry__ as ge
    from oracle import oracle as orc
    dist = _init(rank, world, port, "gloo")
    g = ge.load_package()
    dims, n_rows, gB, iters, step, mu = [12, 9, 7, 4], 64, 8, 5, 0.05, 0.9
    rng = np.random.RandomState(7)
    x = rng.rand(n_rows, dims[0])
    lab = rng.randint(0, dims[-1], n_rows)
    t = np.zeros((n_rows, dims[-1]))
    t[np.arange(n_rows), lab] = 1.0
    idx = g.Sampler(n_rows, seed=1).draw(iters, gB)          # identical on every rank
    mine = g.dp.shard_index_lists(idx, rank, world)
    # the slices tile the global batch exactly
    gathered = [None] * world
    dist.all_gather_object(gathered, mine.tolist())
    assert np.array_equal(np.concatenate([np.asarray(a) for a in gathered], axis=1), idx)

    net = orc.OracleNet(dims)
    w, v = net.get_weights(), np.zeros(net.get_weights().size)
    for it in range(iters):
        net.set_weights(w)
        gsum = np.zeros_like(w)
        for r in mine[it]:
            gsum += net.weight_gradient(x[r], t[r])
        tg = torch.from_numpy(gsum)
        dist.all_reduce(tg)                                    # the cross-GPU batch sum
        v = step * tg.numpy() / gB + mu * v                    # S:333, divide by the GLOBAL batch
        w = w - v
    if rank == 0:
        np.save(os.path.join(out_dir, "dp_w.npy"), w)
        np.save(os.path.join(out_dir, "dp_idx.npy"), idx)
    dist.barrier()
    dist.destroy_process_group()


class _RecordingNet:
    """stands in for a device net: records the order of the b2nn_comm_* calls init_comm makes"""

    def __init__(self, rank, multicast):
        self.rank, self.multicast, self.calls = rank, multicast, []

    def commInit(self, uid, rank, world):
        self.calls.append(("init", bytes(uid), rank, world))

    def commIpcExport(self):
        self.calls.append(("ipc_export",))
        return bytes([self.rank]) * 64

    def commIpcImport(self, handles):
        self.calls.append(("ipc_import", [h[0] for h in handles]))

    def commNvlsExport(self):
        self.calls.append(("nvls_export",))
        return (1000 + self.rank, 50 + self.rank) if self.multicast else None

    def commNvlsImport(self, pairs):
        self.calls.append(("nvls_import", [tuple(p) for p in pairs]))

    def commNvlsTeam(self, pid_fd=None):
        self.calls.append(("nvls_team", pid_fd))
        return (1000, 99)

    def commNvlsBind(self):
        self.calls.append(("nvls_bind",))


def gloo_init_comm_worker(rank, world, port, out_dir):
    """dp.init_comm's three set-ups over gloo with a recording net: what travels between the ranks and in which
    order the library is called (the device side of the same sequence runs in test_world2_matches_single_gpu)"""
    import pickle
    from unittest import mock
    import __graft_entry__ as ge
    dist = _init(rank, world, port, "gloo")
    g = ge.load_package()
    out = {}
    with mock.patch.object(g.nets, "commUniqueId", lambda: b"u" * 128):
        for mode, multicast in [("nccl", True), ("p2p", True), ("nvls", True), ("nvls", False)]:
            os.environ["B2NN_DP"] = mode
            net = _RecordingNet(rank, multicast)
            g.dp.init_comm(net, dist)
            out[(mode, multicast)] = net.calls
        # a TF32 net takes the same exchange set-ups as a BF16 net
        for mode in ("p2p", "nvls"):
            os.environ["B2NN_DP"] = mode
            net = _RecordingNet(rank, True)
            net.precision = g.capi.PREC_TF32
            g.dp.init_comm(net, dist)
            out[("tf32", mode)] = net.calls
        # multicast support on rank 0 only: every rank must raise
        os.environ["B2NN_DP"] = "nvls"
        net = _RecordingNet(rank, rank == 0)
        try:
            g.dp.init_comm(net, dist)
            out["mixed"] = "no exception"
        except RuntimeError as e:
            out["mixed"] = str(e)
        # a stage failing on ONE rank must raise on EVERY rank (nobody is left waiting in a collective)
        os.environ["B2NN_DP"] = "nvls"
        net = _RecordingNet(rank, True)
        if rank == 1:
            def boom(pairs):
                raise ValueError("cuMemImportFromShareableHandle -> 999")
            net.commNvlsImport = boom
        try:
            g.dp.init_comm(net, dist)
            out["failure"] = "no exception"
        except RuntimeError as e:
            out["failure"] = str(e)
    os.environ.pop("B2NN_DP")
    with open(os.path.join(out_dir, "calls_%d.pkl" % rank), "wb") as f:
        pickle.dump(out, f)
    dist.barrier()
    dist.destroy_process_group()


def nccl_worker(rank, world, port, out_dir, precision_name="bf16", extra_flags=0):
    """BF16 (or TF32) nets, one per GPU, stepping on their slice of each global batch with the per-layer NCCL
    all-reduce inside libb2nn.so; rank 0 also steps an un-sharded net on the whole global batch."""
    import torch
    import __graft_entry__ as ge
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    g = ge.load_package()
    rows_per_rank = int(os.environ.get("B2NN_TEST_ROWS_PER_RANK", "256"))
    dims, gB, iters, step, mu = [256, 384, 128, 40], rows_per_rank * world, 6, 0.05, 0.9
    prec = g.capi.PREC_TF32 if precision_name == "tf32" else g.capi.PREC_BF16
    flags = g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX | extra_flags
    rng = np.random.RandomState(11)
    n_rows = 4 * gB
    x = rng.rand(n_rows, dims[0]).astype(np.float32)
    lab = rng.randint(0, dims[-1], n_rows).astype(np.int32)
    idx = g.Sampler(n_rows, seed=1).draw(iters, gB)
    net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=prec,
                                         flags=flags, device=rank)
    g.dp.init_comm(net, dist)
    mine = g.dp.shard_index_lists(idx, rank, world)
    for it in range(iters):
        net.gradientStepLabeled(x[mine[it]], lab[mine[it]], step, mu)
    w = net.getWeights()
    # every replica must hold the same weights bit for bit (same all-reduced gradient, same update)
    tw = torch.from_numpy(w).cuda()
    ref = tw.clone()
    dist.broadcast(ref, src=0)
    same = bool(torch.equal(tw, ref))
    flag = torch.tensor([1 if same else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        single = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=prec,
                                                flags=flags & ~g.capi.FLAG_BF16_WIRE, device=0)
        w0 = single.getWeights()
        for it in range(iters):
            single.gradientStepLabeled(x[idx[it]], lab[idx[it]], step, mu)
        w_single = single.getWeights()
        if os.environ.get("B2NN_TEST_DIAG"):   # a second comparator, stepped with the per-step loss read back
            again = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=prec,
                                                   flags=flags & ~g.capi.FLAG_BF16_WIRE, device=0)
            ls = [again.gradientStepLabeled(x[idx[it]], lab[idx[it]], step, mu, returnLoss=True) for it in range(iters)]
            sys.stderr.write("[dp worker] second comparator: losses %s |dW| %.4g launches %d / %d\n" % (
                ["%.4f" % l for l in ls], np.linalg.norm(again.getWeights() - w0), single.kernelLaunches(),
                again.kernelLaunches()))
        sys.stderr.write("[dp worker] world %d %s %s: |dW_dp| %.4g |dW_single| %.4g finite %s %s\n" % (
            world, precision_name, net.commMode(), np.linalg.norm(w - w0), np.linalg.norm(w_single - w0),
            np.isfinite(w).all(), np.isfinite(w_single).all()))
        np.savez(os.path.join(out_dir, "nccl.npz"), w_dp=w, w_single=w_single, w0=w0,
                 replicas_equal=int(flag.item()), launches=net.kernelLaunches(), mode=net.commMode())
    dist.barrier()
    dist.destroy_process_group()


def timeout_worker(rank, world, port, out_dir):
    """rank 1 joins the group but never takes the step: rank 0's exchange kernel must give up after B2NN_DP_TIMEOUT_MS
    and the call must come back with an error instead of hanging"""
    import time
    import torch
    import __graft_entry__ as ge
    os.environ["B2NN_DP_TIMEOUT_MS"] = "300"
    os.environ["B2NN_DP"] = "p2p"
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    g = ge.load_package()
    dims = [256, 128, 40]
    net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                         flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX, device=rank)
    g.dp.init_comm(net, dist)
    rng = np.random.RandomState(3)
    x = rng.rand(64, dims[0]).astype(np.float32)
    lab = rng.randint(0, dims[-1], 64).astype(np.int32)
    outcome = "no error"
    if rank == 0:
        t0 = time.time()
        try:
            net.gradientStepLabeled(x, lab, 0.05, 0.9)
            net.sync()
        except g.capi.B2nnError as e:
            outcome = "B2nnError %d after %.1f s: %s" % (e.code, time.time() - t0, e)
        with open(os.path.join(out_dir, "timeout.txt"), "w") as f:
            f.write(outcome)
    dist.barrier()      # rank 1 stays alive (its arena mapped) until rank 0 is through
    dist.destroy_process_group()
