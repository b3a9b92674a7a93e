#!/usr/bin/env python
"""bench.py -- headline measurement of the b2nn hot path (train samples/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b2nn|reference] [--workload c4|c5|c2|c1|c3]

One "step" = one mini-batch gradient step (forward, backprop, momentum-SGD update; the reference's
gradientStep, src/SoftmaxCrossEntropyNeuralNet.java:297-346).  Default workload is BASELINE.json's 4096-wide MLP
(configs[3]: synthetic 4096-4096-4096-1000, batch 4096 per GPU, BF16 tcgen05 GEMMs); with N > 1 ranks the global batch
is N x 4096 (weak scaling), each rank runs forward/backward on its 4096 rows and every layer's gradient is exchanged by
ONE fused reduce-scatter + momentum update + all-gather kernel over NVLink (in-switch reduction from 4 GPUs up;
B2NN_DP=nccl selects the NCCL all-reduce + update baseline).  BASELINE configs[4] (8192-wide, 6 matrices, GLOBAL batch
32768 = strong scaling) is measured beside it at every N (`also.c5`), the tiny FP64 configs and the TF32 mode at N = 1.
Prints ONE JSON line on rank 0 (contract: see DESIGN.md "Measurement").

--impl reference times the reference's CPU algorithm for the same workload: the Java trainer cannot run here (no JVM
on this box), so it is the oracle's faithful C restatement, 1 thread like the serial reference, on a bounded sample of
the same workload.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: dims, batch (per GPU for weak scaling, GLOBAL for strong scaling), precision, step size
    "c4": dict(dims=[4096, 4096, 4096, 1000], batch=4096, prec="bf16", step=0.01, scaling="weak",
               desc="synthetic 4096-4096-4096-1000 MLP, batch 4096 per GPU, BF16 tcgen05 GEMMs (BASELINE configs[3])"),
    "c5": dict(dims=[8192] * 6 + [1000], batch=32768, prec="bf16", step=0.01, scaling="strong",
               desc="synthetic 8192-wide 6-matrix MLP (8192x6 -> 1000), GLOBAL batch 32768 split over the GPUs "
                    "(BASELINE configs[4])"),
    "c2": dict(dims=[784, 300, 100, 10], batch=16, prec="f64", step=0.0125, scaling="replicas",
               desc="784-300-100-10 Softmax/CE, batch 16, FP64 parity mode (BASELINE configs[1])"),
    "c1": dict(dims=[784, 100, 50, 10], batch=8, prec="f64", step=0.01, scaling="replicas",
               desc="784-100-50-10 SoftmaxCrossEntropyNeuralNet, batch 8, FP64 parity mode (BASELINE configs[0], "
                    "logs/trainLog.csv:3)"),
    "c3": dict(dims=[784, 100, 50, 10], batch=8, prec="f64", step=0.01, general=True, scaling="replicas",
               desc="GeneralNeuralNet 784-100-50-10, sigmoid inner / outer, half squared error, batch 8, FP64 parity "
                    "mode (BASELINE configs[2])"),
}
if os.environ.get("B2NN_BENCH_C5_ROWS"):   # measurement aid: the per-GPU shape of the 8-GPU C5 step (4096 rows) on one GPU
    WORKLOADS["c5"]["batch"] = int(os.environ["B2NN_BENCH_C5_ROWS"])
    WORKLOADS["c5"]["desc"] += " [global batch overridden to %d by B2NN_BENCH_C5_ROWS]" % WORKLOADS["c5"]["batch"]
CPU_SAMPLE_ROWS = {"c4": 4, "c5": 2, "c2": 16, "c1": 8, "c3": 8}
CPU_SAMPLE_STEPS = {"c4": (10, 1), "c5": (3, 1), "c2": (500, 20), "c1": (3000, 50), "c3": (1500, 50)}  # ~10-15 s each

_REAL_STDOUT = None


def emit(obj):
    f = _REAL_STDOUT or sys.stdout
    f.write(json.dumps(obj) + "\n")
    f.flush()


def n_weights(dims):
    return sum(dims[i] * dims[i + 1] for i in range(len(dims) - 1))


def flops_per_sample(dims):
    # 2 (forward) + 2 (dW) + 2 (dX) per weight, no dX for the first matrix
    return 6 * n_weights(dims) - 2 * dims[0] * dims[1]


def measured_peaks(tf32=False):
    """roofline denominators: MEASURED_PEAKS.json (driver-written cuBLAS BF16 rates: burst = a kernel timed alone,
    sustained = seconds-long loop under the power cap); profiles/tf32_peaks.json holds the same two numbers for
    torch.matmul FP32 with allow_tf32, measured by scratch/tf32_peak.py on this pool's B200"""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        pk = dict(burst=float(d["bf16_tflops"]), sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                  hbm=float(d["hbm_gbs"]), src="measured (MEASURED_PEAKS.json: cuBLAS bf16 8192^3)")
    else:
        pk = dict(burst=1590.0, sustained=1400.0, hbm=6650.0, src="fallback (B200_PROFILING.md)")
    if tf32:
        tp = os.path.join(ROOT, "profiles", "tf32_peaks.json")
        if os.path.exists(tp):
            with open(tp) as f:
                d = json.load(f)
            pk = dict(pk, burst=float(d["tf32_tflops"]), sustained=float(d["tf32_tflops_sustained"]),
                      src="measured (profiles/tf32_peaks.json: torch.matmul fp32 allow_tf32 8192^3, same method as "
                          "MEASURED_PEAKS.json)")
        else:
            pk = dict(pk, burst=1100.0, sustained=1100.0, src="nominal dense TF32 peak (B200_PROFILING.md): no measured "
                                                              "TF32 peak on file")
    return pk


class ClockSampler:
    """SM clock and clock-event reasons of one GPU while the timed regions run, sampled by a separate
    `nvidia-smi -lms` PROCESS (the profiling recipe's clocks line): polling NVML from a thread of this process
    takes driver locks that stall kernel launches and showed up as 3x outliers of the wall-clock e2e number."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        self.index, self.proc, self.t0 = index, None, None
        self.windows = []   # (t0, t1, tag)
        self.rows = None

    def start(self):
        import subprocess
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=timestamp,clocks.sm,clocks.max.sm,"
                 "clocks_event_reasons.active", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def begin(self):
        self.t0 = time.time()

    def end(self, tag="main"):
        if self.t0 is not None:
            self.windows.append((self.t0, time.time(), tag))
            self.t0 = None

    def _collect(self):
        import datetime
        if self.rows is not None:
            return
        self.rows = []
        if self.proc is None:
            return
        time.sleep(0.05)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        for line in out.splitlines():
            f = [q.strip() for q in line.split(",")]
            if len(f) < 4:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                self.rows.append((ts, int(f[1]), int(f[2]), int(f[3], 16)))
            except Exception:
                continue

    def summary(self, tags=None):
        """median SM clock and the reasons seen inside the windows whose tag is in `tags` (all windows when None);
        ends the sampling process on first use"""
        self._collect()
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["clock_sampling_unavailable"]}
        clocks, reasons, mx = [], set(), None
        for ts, mhz, mxv, mask in self.rows:
            if not any(a - 0.02 <= ts <= b + 0.02 for a, b, t in self.windows if tags is None or t in tags):
                continue
            clocks.append(mhz)
            mx = mxv
            for bit, name in self.REASONS.items():
                if mask & bit:
                    reasons.add(name)
        clocks.sort()
        return {"sm_mhz": (clocks[len(clocks) // 2] if clocks else None), "sm_max_mhz": mx, "samples": len(clocks),
                "samples_total": len(self.rows), "reasons": sorted(reasons),
                "how": "nvidia-smi -lms 20 subprocess, samples inside the timed regions"}


def synth_data(rows, d0, n_classes, seed, dtype):
    """x ~ U[0,1) (pixel-like), labels uniform; seeded.  FP64 rows are made the way the
    reference makes MNIST rows -- (byte & 0xff) / 255.0, MNISTTrainer.java:97 -- with MNIST's ~80 % zeros."""
    rng = np.random.default_rng(seed)
    labels = rng.integers(0, n_classes, rows).astype(np.int32)
    if dtype == np.float64:
        raw = rng.integers(0, 256, (rows, d0), dtype=np.uint8)
        raw[rng.random((rows, d0), dtype=np.float32) < 0.8] = 0
        return raw.astype(np.float64) / 255.0, labels
    x = rng.random((rows, d0), dtype=np.float32)
    return x.astype(dtype, copy=False), labels


def cpu_reference_run(workload, steps, warmup, threads=1, rows=None):
    """the oracle's restatement of gradientStep on a bounded sample: faithful single-thread loops (threads = 1, the timed
    CPU baseline), or -- threads > 1, labelled "not the reference" -- the same per-sample arithmetic without the per-sample
    gradient allocation, the samples of a batch spread over POSIX threads (bit-identical results, tests/test_oracle.py)"""
    from oracle import oracle as orc
    wl = WORKLOADS[workload]
    dims, rows = wl["dims"], rows or CPU_SAMPLE_ROWS[workload]
    faithful = threads == 1
    orc.set_threads(threads)
    if wl.get("general"):
        net = orc.OracleNet(dims, kind=orc.NET_GENERAL, inner_act="sigmoid", inner_param=0.0, last_act="sigmoid")
    else:
        net = orc.OracleNet(dims)
    if wl["prec"] != "f64":  # same fan-in-scaled init as the GPU arm so exp() stays finite
        w = net.get_weights()
        off = 0
        for l in range(len(dims) - 1):
            n = dims[l] * dims[l + 1]
            w[off:off + n] *= 2.0 * np.sqrt(6.0 / dims[l])
            off += n
        net.set_weights(w)
    x, labels = synth_data(rows * (steps + warmup), dims[0], dims[-1], 1, np.float64)
    t = np.zeros((x.shape[0], dims[-1]))
    t[np.arange(x.shape[0]), labels] = 1.0
    for s in range(warmup):
        net.gradient_step(x[s * rows:(s + 1) * rows], t[s * rows:(s + 1) * rows], wl["step"], 0.9, faithful=faithful)
    t0 = time.perf_counter()
    for s in range(warmup, warmup + steps):
        net.gradient_step(x[s * rows:(s + 1) * rows], t[s * rows:(s + 1) * rows], wl["step"], 0.9, faithful=faithful)
    dt = time.perf_counter() - t0
    orc.set_threads(1)
    return rows * steps / dt, dt, rows


def cpu_baseline_block(workload):
    cs, cw = CPU_SAMPLE_STEPS[workload]
    val, dt, r = cpu_reference_run(workload, cs, cw)
    blk = {"value": val, "unit": "samples/s", "cores": 1, "kind": "port",
           "sample": "%d gradient steps of %d samples of the same net by the oracle's faithful loops (%.1f s)" % (cs, r, dt)}
    if WORKLOADS[workload]["prec"] != "f64":
        # second line (optional): best-effort CPU on every host core -- NOT the reference, which is serial
        T = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        if T > 1:
            v2, dt2, r2 = cpu_reference_run(workload, 2, 1, threads=T, rows=4 * T)
            blk["all_cores"] = {"value": v2, "unit": "samples/s", "cores": T,
                                "kind": "port, not the reference (which is serial): the same per-sample arithmetic, the "
                                        "samples of a batch spread over POSIX threads, no per-sample gradient allocation",
                                "sample": "2 gradient steps of %d samples (%.1f s)" % (r2, dt2)}
    return blk


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    val, dt, rows = cpu_reference_run(args.workload, args.steps, args.warmup)
    sample = ("%d steps of %d samples each of the same net (faithful per-sample loops, per-sample gradient "
              "allocation, batch sum, momentum update); cost is linear in samples" % (args.steps, rows))
    out = {
        "impl": "reference", "metric": "train_samples_per_sec", "value": val, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong" if wl["scaling"] == "strong" else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "dims": wl["dims"], "batch_per_step": rows},
        "cpu_baseline": {"value": val, "unit": "samples/s", "cores": 1, "kind": "port", "sample": sample,
                         "note": "the Java reference cannot run on this box (no JVM); oracle = C restatement, "
                                 "single thread like the serial reference (README.md:13)"},
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if wl["prec"] != "f64":
        # beside it, never instead of it: the threaded port on every host core (not the reference, which is serial)
        T = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        if T > 1:
            v2, dt2, r2 = cpu_reference_run(args.workload, 2, 1, threads=T, rows=4 * T)
            out["cpu_baseline"]["all_cores"] = {
                "value": v2, "unit": "samples/s", "cores": T, "sample": "2 gradient steps of %d samples (%.1f s)" % (r2, dt2),
                "kind": "port, not the reference (which is serial): the same per-sample arithmetic, the samples of a batch "
                        "spread over POSIX threads, no per-sample gradient allocation"}
    emit(out)


class Ctx:
    """what every measurement needs: the package, torch, the process group, this rank, the clock sampler"""

    def __init__(self, g, torch, dist, rank, world, local_rank, sampler):
        self.g, self.torch, self.dist = g, torch, dist
        self.rank, self.world, self.local_rank, self.sampler = rank, world, local_rank, sampler
        self.stream = torch.cuda.current_stream()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def events(self):
        return self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)


def train_call(g, net, ds, idx, B, step, want_loss=False):
    loss = np.empty(idx.shape[0]) if want_loss else None
    g.capi.check(g.capi.lib().b2nn_train_indexed(
        net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), idx.shape[0], B, step, 0.9,
        loss.ctypes.data_as(C.POINTER(C.c_double)) if want_loss else None))
    return loss


def dp_check(ctx, net, make_single, dims, B, step):
    """Data-parallel correctness inside the driver's own run: 3 steps with every rank on its B rows of a global batch of
    world x B rows, against the same 3 steps of ONE net on rank 0 over the whole global batch (the reference's batch
    sum S:305-322 divided by batch.size() S:333).  rel = |dW_dp - dW_1| / |dW_1| (only the FP32 order of the batch
    sum differs; bound 2e-3 as in tests/test_dp.py); replicas_equal = every rank holds bit-identical weights."""
    g, dist, rank, world = ctx.g, ctx.dist, ctx.rank, ctx.world
    rng = np.random.default_rng(99)
    n = world * B
    x = rng.random((n, dims[0]), dtype=np.float32)
    lab = rng.integers(0, dims[-1], n).astype(np.int32)
    w0 = net.getWeights() if rank == 0 else None
    mine = slice(rank * B, (rank + 1) * B)
    for _ in range(3):
        net.gradientStepLabeled(x[mine], lab[mine], step, 0.9)
    net.sync()
    ctx.barrier()
    w_dp = net.getWeights()
    digest = hashlib.sha1(w_dp.astype(np.float32).tobytes()).hexdigest()
    digests = [None] * world
    dist.all_gather_object(digests, digest)
    res = [None]
    if rank == 0:
        single = make_single()
        for _ in range(3):
            single.gradientStepLabeled(x, lab, step, 0.9)
        w1 = single.getWeights()
        single.close()
        du_dp, du_1 = w_dp - w0, w1 - w0
        res[0] = {"rel": float(np.linalg.norm(du_dp - du_1) / np.linalg.norm(du_1)),
                  "replicas_equal": all(d == digests[0] for d in digests), "exchange": net.commMode(), "steps": 3,
                  "global_batch": n, "bound": 2e-3,
                  "what": "3 data-parallel steps vs the same 3 steps of one net on rank 0 over the global batch"}
    dist.broadcast_object_list(res, src=0)
    ctx.barrier()
    return res[0]


def measure_wide(ctx, name, tf32, K, W, want_e2e=True, want_sustained=False, want_dp_check=True, e2e_rows="u8"):
    """one tensor-core workload: device-resident timed region, serialised roofline pass, optional end-to-end leg from
    pinned host rows, optional sustained (>= 2 s back-to-back) leg"""
    g, torch, dist, rank, world, dev = ctx.g, ctx.torch, ctx.dist, ctx.rank, ctx.world, ctx.local_rank
    wl = WORKLOADS[name]
    dims, step = wl["dims"], wl["step"]
    strong = wl["scaling"] == "strong"
    B = wl["batch"] // world if strong else wl["batch"]        # rows per GPU per step
    gB = B * world
    prec_name = "tf32" if tf32 else "bf16"
    # --wire bf16 (opt-in, B2NN_FLAG_BF16_WIRE): per-GPU gradients exchanged as BF16; dp_check reports what it costs
    wire_flag = g.capi.FLAG_BF16_WIRE if (os.environ.get("B2NN_BENCH_WIRE") == "bf16" and not tf32) else 0

    def make_net():
        return g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME,
                                              precision=g.capi.PREC_TF32 if tf32 else g.capi.PREC_BF16,
                                              flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX | wire_flag,
                                              device=dev)

    net = make_net()
    if world > 1:
        try:
            g.dp.init_comm(net, dist, rank, world)
        except RuntimeError as e:
            # the in-switch exchange could not be set up (raised on every rank alike, dp._agree): its arena has
            # already moved, so start over with plain peer memory
            sys.stderr.write("[bench] %s -- falling back to B2NN_DP=p2p\n" % e)
            net.close()
            os.environ["B2NN_DP"] = "p2p"
            net = make_net()
            g.dp.init_comm(net, dist, rank, world)
    net.setStream(ctx.stream.cuda_stream)
    res = {"workload": name + ": " + (wl["desc"].replace("BF16", "TF32") if tf32 else wl["desc"]), "dims": dims,
           "dtype": prec_name, "batch_per_gpu": B, "global_batch": gB, "scaling": wl["scaling"],
           "exchange": net.commMode(), "gradient_wire": "bf16" if (wire_flag and world > 1) else "fp32"}

    if world > 1 and want_dp_check:
        res["dp_check"] = dp_check(ctx, net, make_net, dims, B, step)

    rows = (16 if name == "c4" else 2) * B
    x, labels = synth_data(rows, dims[0], dims[-1], 1 + rank, np.float32)
    ds_dev = g.Dataset(net, x, labels=labels, location=g.capi.DATA_DEVICE)
    smp = g.Sampler(rows, seed=1)
    idx_warm = smp.draw(max(W, 1), B)
    idx_time = smp.draw(K, B)

    # ---- device-resident timed region -----------------------------------------------------
    train_call(g, net, ds_dev, idx_warm, B, step)
    ctx.barrier()
    l0 = net.kernelLaunches()
    ev0, ev1 = ctx.events()
    ctx.sampler.begin()
    ev0.record(ctx.stream)
    train_call(g, net, ds_dev, idx_time, B, step)
    ev1.record(ctx.stream)
    ctx.barrier()
    ctx.sampler.end(name)
    ms = ctx.max_over_ranks(ev0.elapsed_time(ev1))
    res["gpu_launches"] = net.kernelLaunches() - l0
    res["ms_per_step"] = ms / K
    res["value"] = gB * K / (ms * 1e-3)
    res["steps"] = K
    fl_step_gpu = float(B * flops_per_sample(dims))
    res["step_tflops_per_gpu"] = fl_step_gpu * K / (ms * 1e-3) / 1e12

    # ---- roofline passes: the same K steps under CUDA events on the launching stream ---------------------------------
    # pass A (b2nn_profile_enable 2): one event pair around each run of consecutive contraction launches (the forward
    #   GEMMs, the backward GEMMs): roofline.achieved.  pass B (1): one pair around EVERY launch, which times each kernel
    #   class alone (non_gemm_us) but idles the stream for several us per event, part of it inside the next window --
    #   its per-launch average is kept beside pass A's as avg_launch_ms_event_per_launch.
    def roofline_pass(mode):
        net.profileEnable(mode)
        ctx.sampler.begin()
        e2, e3 = ctx.events()
        e2.record(ctx.stream)
        train_call(g, net, ds_dev, idx_time, B, step)
        e3.record(ctx.stream)
        read, by_class = net.profileRead(), net.profileReadClasses()
        ctx.sampler.end(name + ":roofline")
        net.profileEnable(False)
        return read, by_class, e2.elapsed_time(e3) / K

    (gemm_ms, gemm_n, gemm_flops), classes_a, prof_step_ms = roofline_pass(2)
    (gemm_ms_b, gemm_n_b, _), classes, prof_step_ms_b = roofline_pass(1)
    # pass C (3): no events, nothing serialised -- the contraction kernels stamp %globaltimer themselves (first CTA in, last
    #   CTA out); the union of those intervals is the time the step really spends inside contractions
    net.profileEnable(3)
    e4, e5 = ctx.events()
    e4.record(ctx.stream)
    train_call(g, net, ds_dev, idx_time, B, step)
    e5.record(ctx.stream)
    busy_ms, stamp_sum_ms, stamp_n, stamp_flops = net.profileReadStamps()
    stamp_mhz = net.profileStampClockMHz()
    net.profileEnable(False)
    stamp_step_ms = e4.elapsed_time(e5) / K
    pk = measured_peaks(tf32)
    achieved = gemm_flops / (gemm_ms * 1e-3) / 1e12
    # a timed region shorter than ~1 s runs at boost clocks (burst peak); longer ones meet the 1 kW cap (sustained peak)
    region_s = ms * 1e-3
    which = "burst" if region_s < 1.0 else "sustained"
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        with open(tp) as f:
            traffic = json.load(f).get(name + ("_tf32" if tf32 else ""))
    gemm_us = gemm_ms / K * 1e3
    res["roofline"] = {
        "bound": "tensor", "kernel": "gemm_%s_pair_kernel (tcgen05 cta_group::2, all epilogues)" % prec_name,
        "achieved": achieved, "peak": pk[which], "unit": "TFLOP/s", "frac": achieved / pk[which],
        "peak_kind": "%s (timed region %.3f s)" % (which, region_s), "frac_burst": achieved / pk["burst"],
        "frac_sustained": achieved / pk["sustained"], "peak_burst": pk["burst"], "peak_sustained": pk["sustained"],
        "peak_source": pk["src"], "traffic": traffic, "launches": gemm_n, "avg_launch_ms": gemm_ms / max(gemm_n, 1),
        "windows": classes_a["contraction"][1], "how": "CUDA events on the launching stream, one pair per run of "
        "consecutive contraction launches (forward pass; backward GEMMs), same K steps as the timed region",
        "avg_launch_ms_event_per_launch": gemm_ms_b / max(gemm_n_b, 1),
        # the same K steps with no events at all (production overlap): in-kernel %globaltimer stamps per launch
        "in_kernel_timer": {"contraction_busy_ms_per_step": busy_ms / K, "avg_launch_ms": stamp_sum_ms / max(stamp_n, 1),
                            "launches": stamp_n, "achieved": stamp_flops / max(busy_ms, 1e-9) / 1e9,
                            "frac_burst": stamp_flops / max(busy_ms, 1e-9) / 1e9 / pk["burst"],
                            "frac_sustained": stamp_flops / max(busy_ms, 1e-9) / 1e9 / pk["sustained"],
                            "step_ms": stamp_step_ms,
                            # cycles / ns of CTA 0 of every launch: the SM clock the timed steps really ran at (nvidia-smi
                            # samples every ~20 ms and misses what the power cap does inside a 36 ms region)
                            "sm_mhz_effective": stamp_mhz,
                            "how": "b2nn_profile_enable(3): every contraction launch records %globaltimer at its first CTA's "
                                   "start and its last CTA's end; busy = union of the intervals, nothing serialised"},
        "flops_per_step": fl_step_gpu, "gemm_share_of_step": (gemm_ms / K) / prof_step_ms,
        "serialised_step_ms": prof_step_ms, "serialised_step_ms_event_per_launch": prof_step_ms_b, "step_frac_burst": res["step_tflops_per_gpu"] / pk["burst"],
        "step_frac_sustained": res["step_tflops_per_gpu"] / pk["sustained"],
        # where the rest of the step goes: each kernel class timed alone in the serialised pass (us per step), and
        # what the overlapped step spends beyond its contractions
        "non_gemm_us": {"output_delta": classes["output"][0] / K * 1e3,
                        "update_or_exchange_side_stream": classes["update"][0] / K * 1e3,
                        "input_prologue_copy_stream": classes["prologue"][0] / K * 1e3,
                        "contractions_alone": gemm_us,
                        "step_minus_contractions_alone": res["ms_per_step"] * 1e3 - gemm_us},
    }

    # ---- end to end: rows stay in pinned host memory, every step's batch crosses PCIe, per-step loss read back -------
    if want_e2e:
        def e2e_leg(kind, sample_clocks):
            """K steps through b2nn_train_indexed over a B2NN_DATA_HOST_PINNED dataset whose rows are `kind`; wall clock around
            the call, max over ranks"""
            if kind == "u8":
                x_host = np.minimum(np.floor(x * 256.0), 255.0).astype(np.uint8)   # IDX pixel bytes, (raw & 0xff) / 255.0 (M:97)
            elif kind == "f32":
                x_host = x                                                          # FP32 rows: the operand container of TF32 nets
            else:
                u = x.view(np.uint32)                                               # BF16 bits, rounded to nearest even once at staging
                x_host = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
            ds_h = g.Dataset(net, x_host, labels=labels, location=g.capi.DATA_HOST_PINNED)
            # warm-up = one untimed pass over the SAME index lists: the timed pass then touches no pinned page for the first
            # time.  With 3 warm-up steps over other rows, the FIRST host-fed leg of a process came out at 1.2-2.4 ms/step
            # in ~1 run of 8, whichever row type it carried (the second leg, and 19 isolated runs that repeated their index
            # lists, never did): a first-touch cost of freshly pinned guest pages on the device side of the VM, not a
            # property of the pipeline.
            # Then a quarter of a second of idle and 3 warm-up steps, so that the timed pass starts from the same power state
            # as the device-resident timed region (a short warm-up after a host phase): straight behind the 50-step warm-up
            # pass the leg runs 7 % slower, like the 200-step regions of profiles/r02_notes.md (the power cap, not the feed).
            train_call(g, net, ds_h, idx_time, B, step, want_loss=True)
            torch.cuda.synchronize()
            time.sleep(0.25)
            train_call(g, net, ds_h, idx_warm[:min(3, idx_warm.shape[0])], B, step, want_loss=True)
            ctx.barrier()
            if sample_clocks:
                ctx.sampler.begin()
            t0 = time.perf_counter()
            losses = train_call(g, net, ds_h, idx_time, B, step, want_loss=True)
            torch.cuda.synchronize()
            sec = ctx.max_over_ranks(time.perf_counter() - t0)
            if sample_clocks:
                ctx.sampler.end(name + ":e2e")
            ctx.barrier()
            ds_h.close()
            return {"value": gB * K / sec, "unit": "samples/s", "ms_per_step": sec / K * 1e3,
                    "h2d_bytes_per_step": B * dims[0] * x_host.dtype.itemsize + B * 4, "last_loss": float(losses[-1]),
                    "rows": kind}
        # Headline leg: the rows in the reference's own storage format -- IDX pixel bytes, read as (raw & 0xff) / 255.0
        # (MNISTTrainer.java:97) and converted by the input prologue kernel.  The BF16-row leg moves twice the bytes
        # (33.6 MB per 0.71 ms step: its copy is busy 0.72 of every 0.80 ms, ~47 GB/s of the PCIe 5 x16 link) and is
        # reported beside the headline.  TF32 nets read FP32 rows.
        first = "f32" if tf32 else e2e_rows
        res["e2e"] = e2e_leg(first, True)
        res["e2e"]["d2h_bytes_per_step"] = 8
        res["e2e"]["path"] = ("b2nn_train_indexed over a B2NN_DATA_HOST_PINNED dataset: every step's batch rows are "
                              "gathered over PCIe from pinned host memory by the input prologue kernel on a copy stream "
                              "(double-buffered under the previous step's GEMMs); per-step loss read back")
        if not tf32 and name == "c4":
            other = "bf16" if first == "u8" else "u8"
            leg = e2e_leg(other, False)
            res["e2e"]["bf16_rows" if other == "bf16" else "pixel_byte_rows"] = {
                k: leg[k] for k in ("value", "unit", "ms_per_step", "h2d_bytes_per_step")}

    # ---- sustained leg (last: it heats the part up): the same call back to back for >= 2 s, power-capped clocks ------
    if want_sustained:
        train_call(g, net, ds_dev, idx_time, B, step)
        ctx.barrier()
        ev4, ev5 = ctx.events()
        ctx.sampler.begin()
        t0 = time.perf_counter()
        ev4.record(ctx.stream)
        calls = 0
        while time.perf_counter() - t0 < 2.0 or calls < 2:
            for _ in range(4):
                train_call(g, net, ds_dev, idx_time, B, step)
            calls += 4
            torch.cuda.synchronize()
        ev5.record(ctx.stream)
        ctx.barrier()
        ctx.sampler.end(name + ":sustained")
        sms = ctx.max_over_ranks(ev4.elapsed_time(ev5)) / (calls * K)
        stf = fl_step_gpu / (sms * 1e-3) / 1e12
        res["sustained"] = {"ms_per_step": sms, "steps": calls * K, "value": gB / (sms * 1e-3), "tflops_per_gpu": stf,
                            "step_frac_sustained": stf / pk["sustained"], "clock_tag": name + ":sustained"}
    ds_dev.close()
    net.close()
    return res


def measure_f64(ctx, name, K, W, x=None, labels=None, want_cpu=False):
    """one tiny FP64 parity workload through K1 (f64::fused_train_kernel): device-resident run of K steps in one launch,
    the per-phase cycle counters of a second, profiled run, and the drop-in per-call path (NeuralNet.gradientStep with
    host double[B][784] in, synchronous) in microseconds per call"""
    g, torch, dev = ctx.g, ctx.torch, ctx.local_rank
    wl = WORKLOADS[name]
    dims, B, step = wl["dims"], wl["batch"], wl["step"]
    if wl.get("general"):
        net = g.GeneralNeuralNet(dims, g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME,
                                 g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME, device=dev)
    else:
        net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, device=dev)
    net.setStream(ctx.stream.cuda_stream)
    if x is None:
        x, labels = synth_data(60000, dims[0], dims[-1], 1 + ctx.rank, np.float64)
    ds = g.Dataset(net, x, labels=labels)
    smp = g.Sampler(x.shape[0], seed=1)
    train_call(g, net, ds, smp.draw(max(W, 1), B), B, step)
    ctx.barrier()
    idx = smp.draw(K, B)
    l0 = net.kernelLaunches()
    e0, e1 = ctx.events()
    ctx.sampler.begin()
    e0.record(ctx.stream)
    train_call(g, net, ds, idx, B, step)
    e1.record(ctx.stream)
    ctx.barrier()
    ctx.sampler.end(name)
    ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    launches = net.kernelLaunches() - l0
    res = {"workload": name + ": " + wl["desc"], "dims": dims, "dtype": "f64", "batch_per_gpu": B, "global_batch": B,
           "scaling": "replicas", "value": B * K / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / K,
           "us_per_step": ms / K * 1e3, "steps": K, "gpu_launches": launches}

    # e2e = the same fast path with the per-step loss read back (the dataset is device-resident like the reference's
    # in-memory HashMap; index lists cross PCIe every call)
    ctx.sampler.begin()
    t0 = time.perf_counter()
    losses = train_call(g, net, ds, idx, B, step, want_loss=True)
    torch.cuda.synchronize()
    e2e_s = ctx.max_over_ranks(time.perf_counter() - t0)
    ctx.sampler.end(name + ":e2e")
    # the drop-in call the Java API makes per step (T:70, T:89): NeuralNet.gradientStep with host rows, synchronous
    nb = 300
    xb = np.ascontiguousarray(x[:B * 8].reshape(8, B, dims[0]))
    tb = np.zeros((8, B, dims[-1]))
    for q in range(8):
        tb[q, np.arange(B), labels[q * B:(q + 1) * B]] = 1.0
    dp = C.POINTER(C.c_double)
    xp = [xb[q].ctypes.data_as(dp) for q in range(8)]
    tp = [tb[q].ctypes.data_as(dp) for q in range(8)]
    fn, h = g.capi.lib().b2nn_gradient_step, net._h
    for q in range(20):
        g.capi.check(fn(h, xp[q % 8], tp[q % 8], B, step, 0.9, 0, None))
    t1 = time.perf_counter()
    for q in range(nb):
        fn(h, xp[q % 8], tp[q % 8], B, step, 0.9, 0, None)
    per_call_us = (time.perf_counter() - t1) / nb * 1e6
    res["e2e"] = {"value": B * K / e2e_s, "unit": "samples/s", "h2d_bytes_per_step": B * 4, "d2h_bytes_per_step": 8,
                  "ms_per_step": e2e_s / K * 1e3, "last_loss": float(losses[-1]), "rows": "f64",
                  "path": "b2nn_train_indexed over the device-resident dataMaster; index lists cross PCIe every call, "
                          "per-step loss read back",
                  "gradient_step_call": {"us_per_call": per_call_us, "value": B / (per_call_us * 1e-6), "unit": "samples/s",
                                         "h2d_bytes_per_call": B * (dims[0] + dims[-1]) * 8, "calls": nb,
                                         "path": "b2nn_gradient_step (NeuralNet.gradientStep, S:297): host double rows "
                                                 "and targets in, one K1 step, stream sync before return"}}

    # K2: batched FP64 inference with the `>=` argmax and the per-row loss (testOnTrainingData / validate, M:159-197,
    # T:102-113) over all staged rows, one b2nn_evaluate call
    ds.evaluate()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        hits, mean_loss, _ = ds.evaluate()
    ev_s = (time.perf_counter() - t2) / reps
    res["k2_evaluate"] = {"rows": int(x.shape[0]), "ms_per_call": ev_s * 1e3, "rows_per_s": x.shape[0] / ev_s,
                          "fp64_gflops": 2.0 * n_weights(dims) * x.shape[0] / ev_s / 1e9,
                          "kernels": "f64::gather_rows + gemm_strict (ascending-k, un-fused multiply-add, like the JVM) per "
                                     "layer + output_rows + reduce_rows, chunks of 8192 rows; wall clock incl. the sync"}

    # K1 phase counters (cycles per step on CTA 0): the latency chain this kernel is bound by
    net.k1Profile(True)
    kp = min(K, 2000)
    train_call(g, net, ds, idx[:kp], B, step)
    net.sync()
    phases, nsteps = net.k1Profile(False)
    total = sum(phases.values())
    P = n_weights(dims)
    fl = float(B * flops_per_sample(dims))
    res["roofline"] = {
        "bound": "latency", "kernel": "f64::fused_train_kernel (K1: one persistent cooperative launch for the whole run)",
        "achieved": ms / K * 1e3, "unit": "us/step", "peak": None, "frac": None, "traffic": None,
        "why_no_peak": "the state (%.1f MB of FP64 weights + velocity) never leaves the chip and a step is %.1f MFLOP: "
                       "neither HBM nor the tensor pipe bounds it, the chain of dependent phases and its grid barriers does"
                       % (P * 16 / 1e6, fl / 1e6),
        "phase_cycles_per_step": phases, "cycles_per_step": total, "profiled_steps": nsteps,
        "barrier_share": (phases.get("barrier1", 0) + phases.get("barrier2", 0)) / total if total else None,
        "fp64_gflops": fl / (ms / K * 1e-3) / 1e9, "flops_per_step": fl,
        "l2_resident_bytes_per_step": 4 * P * 8 + B * (dims[0] + dims[-1]) * 8}
    if want_cpu:
        res["cpu_baseline"] = cpu_baseline_block(name)
    ds.close()
    net.close()
    return res


def run_b2nn(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    g = ge.load_package()
    wl = WORKLOADS[args.workload]
    K, W = args.steps, args.warmup
    wide = wl["prec"] == "bf16"
    tf32 = wide and args.precision == "tf32"
    sampler = ClockSampler(local_rank)
    sampler.start()
    ctx = Ctx(g, torch, dist, rank, world, local_rank, sampler)

    also = {}
    if wide:
        main = measure_wide(ctx, args.workload, tf32, K, W, want_e2e=True, want_sustained=not args.no_sustained,
                            e2e_rows=args.e2e_rows)
        if not args.no_also:
            if args.workload != "c5":
                # BASELINE configs[4] at every N: strong scaling of the 8192-wide net over a global batch of 32768
                try:
                    also["c5"] = measure_wide(ctx, "c5", False, 5, 3, want_e2e=False, want_sustained=False,
                                              want_dp_check=False)
                except Exception as e:  # noqa
                    also["c5"] = {"error": repr(e)}
            if world == 1 and args.workload == "c4" and not tf32:
                try:
                    also["c4_tf32"] = measure_wide(ctx, "c4", True, 20, 5, want_e2e=False, want_sustained=False)
                except Exception as e:  # noqa
                    also["c4_tf32"] = {"error": repr(e)}
                try:
                    x64, l64 = synth_data(60000, 784, 10, 1, np.float64)
                    also["c2"] = measure_f64(ctx, "c2", 20000, 200, x64, l64, want_cpu=not args.no_cpu)
                    also["c1"] = measure_f64(ctx, "c1", 20000, 200, x64, l64)
                    also["c3"] = measure_f64(ctx, "c3", 20000, 200, x64, l64)
                except Exception as e:  # noqa
                    also["f64_error"] = repr(e)
    else:
        main = measure_f64(ctx, args.workload, K, W)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline_block(args.workload)

    if rank == 0:
        dims = wl["dims"]
        if wide:
            par = "dp%d" % world
            l2 = ("no flush needed: each step streams %.0f MB of weight/velocity state plus a fresh batch, larger than "
                  "the 126 MB L2" % (n_weights(dims) * 12 / 1e6))
            scaling = wl["scaling"]
        else:
            # the tiny FP64 configs do not shard: every rank is an independent replica with no exchange;
            # `value` is ONE replica's rate (max step time over the replicas), the replica count is stated beside it
            par = "replicas only x%d (no exchange; value = one replica's rate)" % world
            l2 = ("state is L2-resident by design (%.1f MB of weights + velocity); batches cycle through 60000 rows "
                  "(376 MB)" % (n_weights(dims) * 16 / 1e6))
            scaling = "weak"
        if "sustained" in main:
            clk = sampler.summary({main["sustained"].pop("clock_tag")})
            main["sustained"]["sm_mhz"], main["sustained"]["reasons"] = clk.get("sm_mhz"), clk.get("reasons")
        out = {
            "metric": "train_samples_per_sec", "value": main["value"], "unit": "samples/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": main["dtype"], "data": "synthetic",
            "config": {"workload": main["workload"], "dims": dims, "batch_per_gpu": main["batch_per_gpu"],
                       "global_batch": main["global_batch"], "parallelism": par,
                       "exchange": main.get("exchange", "single"), "gradient_wire": main.get("gradient_wire"),
                       "replicas": world if not wide else None, "l2": l2},
            "roofline": main["roofline"], "cpu_baseline": cpu, "e2e": main["e2e"],
            "gpu_launches": main["gpu_launches"],
            "clocks": sampler.summary({args.workload, args.workload + ":roofline", args.workload + ":e2e"}),
        }
        for k in ("sustained", "dp_check", "step_tflops_per_gpu"):
            if k in main:
                out[k] = main[k]
        if also:
            out["also"] = also
        emit(out)
    if world > 1:
        dist.destroy_process_group()


def main():
    # the contract is ONE JSON line on stdout: libraries that chat on fd 1 (NCCL prints its version there) are
    # sent to stderr, and the JSON line goes to the saved descriptor
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b2nn", choices=["b2nn", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--e2e-rows", default="u8", choices=["bf16", "u8"],
                    help="element type of the pinned host rows of the headline e2e leg (u8 = IDX pixel bytes / 255, the "
                         "reference's storage format; the other one is reported beside it)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "tf32"],
                    help="operand type of the tensor-core workloads c4 / c5")
    ap.add_argument("--wire", default="fp32", choices=["fp32", "bf16"],
                    help="data-parallel gradient exchange format (bf16 = B2NN_FLAG_BF16_WIRE, opt-in: see dp_check.rel)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-also", action="store_true", help="skip the side measurements (c5, TF32, the FP64 configs)")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 2 s back-to-back leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.wire == "bf16":
        os.environ["B2NN_BENCH_WIRE"] = "bf16"
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b2nn(args)


if __name__ == "__main__":
    main()
