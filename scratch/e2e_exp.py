import os, sys, time, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as ge
g = ge.load_package()
dims, B, K = [4096, 4096, 4096, 1000], 4096, 50
net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                     flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
rng = np.random.default_rng(1)
rows = 16 * B
x = rng.random((rows, dims[0]), dtype=np.float32)
lab = rng.integers(0, 1000, rows).astype(np.int32)
xb = (x.view(np.uint32) >> 16).astype(np.uint16)
smp = g.Sampler(rows, seed=1)
idx = smp.draw(K, B)
def run(ds, want_loss):
    loss = np.empty(K) if want_loss else None
    g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), K, B, 0.01, 0.9,
                 loss.ctypes.data_as(C.POINTER(C.c_double)) if want_loss else None))
    net.sync()
xu = np.minimum(np.floor(x * 256.0), 255.0).astype(np.uint8)
for name, arr, loc in [("dev_f32", x, g.capi.DATA_DEVICE), ("dev_bf16", xb, g.capi.DATA_DEVICE), ("pin_u8", xu, g.capi.DATA_HOST_PINNED), ("pin_bf16", xb, g.capi.DATA_HOST_PINNED)]:
    ds = g.Dataset(net, arr, labels=lab, location=loc)
    for wl in (False, True):
        run(ds, wl); run(ds, wl)
        t0 = time.perf_counter(); run(ds, wl); dt = time.perf_counter() - t0
        print(os.environ.get("B2NN_PIPELINE", "1"), os.environ.get("B2NN_PROLOGUE_BPS", "-"), name, "loss" if wl else "noloss", "%.3f ms/step" % (dt / K * 1e3), flush=True)
    ds.close()
