"""Where does the sustained (power-capped) step lose against cuBLAS?  Back-to-back for ~2 s each:
  (a) torch.matmul 4096^3 and 8192^3;
  (b) forward only (evaluate) on a 4096x5 net: our GEMM with the activation epilogue, no side kernels;
  (c) training steps of the same net, profiling on (side work serialised) and off."""
import os, subprocess, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as ge
g = ge.load_package()


class Smi:
    def __init__(self):
        self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits",
                                   "-lms", "50", "-i", "0"], stdout=subprocess.PIPE, text=True)
        import threading
        self.rows = []
        self.t = threading.Thread(target=self._rd, daemon=True)
        self.t.start()

    def _rd(self):
        for line in self.p.stdout:
            try:
                a, b = line.split(",")
                self.rows.append((time.perf_counter(), float(a), float(b)))
            except Exception:
                pass

    def window(self, t0, t1):
        r = [x for x in self.rows if t0 + 0.3 <= x[0] <= t1]
        if not r:
            return "no samples"
        return "sm %d MHz, %d W (n=%d)" % (np.median([x[1] for x in r]), np.median([x[2] for x in r]), len(r))


smi = Smi()
time.sleep(0.5)


def timed(name, fn, flops_per_call, secs=2.0):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 0
    while time.perf_counter() - t0 < secs:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    t1 = time.perf_counter()
    ms = e0.elapsed_time(e1)
    print("%-44s %8.3f ms/call %8.1f TFLOP/s   %s" % (name, ms / n, flops_per_call * n / ms / 1e9, smi.window(t0, t1)), flush=True)
    time.sleep(1.0)


for N in (4096, 8192):
    a = torch.randn(N, N, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, N, device="cuda", dtype=torch.bfloat16)
    c = torch.empty(N, N, device="cuda", dtype=torch.bfloat16)
    timed("torch.matmul %d^3" % N, lambda: torch.matmul(a, b, out=c), 2.0 * N ** 3)
del a, b, c

B = 4096
dims = [4096] * 5 + [4096]
net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                     flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
net.setStream(torch.cuda.current_stream().cuda_stream)
rows = 4 * B
rng = np.random.default_rng(1)
x = rng.random((rows, dims[0]), dtype=np.float32)
lab = rng.integers(0, dims[-1], rows).astype(np.int32)
ds = g.Dataset(net, x, labels=lab, location=g.capi.DATA_DEVICE)
L = len(dims) - 1
fwd = 2.0 * B * 4096 * 4096 * L
timed("evaluate (forward only), %d rows" % rows, lambda: ds.evaluate(rows, want_loss=False), fwd * rows / B)

import ctypes as C
idx = np.ascontiguousarray(g.Sampler(rows, seed=1).draw(10, B), dtype=np.int32)


def train10():
    g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), 10, B,
                                                 C.c_double(0.001), C.c_double(0.9), None))
step_flops = fwd + fwd + 2.0 * B * 4096 * 4096 * (L - 1)
for env in ({}, {"prof": 1}):
    if env.get("prof"):
        net.profileEnable(True)
    timed("train 10 steps%s" % (" [profiling, serialised]" if env else ""), train10, step_flops * 10)
    if env.get("prof"):
        ms, n, fl = net.profileRead()
        print("   GEMMs alone: %d launches, %.4f ms avg, %.1f TFLOP/s" % (n, ms / n, fl / ms / 1e9))
        net.profileEnable(False)
smi.p.terminate()
