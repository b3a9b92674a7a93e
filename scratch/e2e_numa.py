import os, sys, time, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as ge
g = ge.load_package()
dims, B, K = [4096, 4096, 4096, 1000], 4096, 30
net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                     flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
rng = np.random.default_rng(1)
rows = 16 * B
x = rng.random((rows, dims[0]), dtype=np.float32)
lab = rng.integers(0, 1000, rows).astype(np.int32)
xb = (x.view(np.uint32) >> 16).astype(np.uint16)
idx = g.Sampler(rows, seed=1).draw(K, B)
ds = g.Dataset(net, xb, labels=lab, location=g.capi.DATA_HOST_PINNED)
def run():
    loss = np.empty(K)
    g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), K, B, 0.01, 0.9,
                 loss.ctypes.data_as(C.POINTER(C.c_double))))
run()
t0 = time.perf_counter(); run(); dt = time.perf_counter() - t0
print(sys.argv[1:], "affinity", sorted(os.sched_getaffinity(0))[:4], "...", "%.3f ms/step" % (dt / K * 1e3), flush=True)
