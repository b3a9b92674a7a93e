"""B2NN_GEMM_STALLS=1: one C4 training step, prints where each GEMM's role warps wait"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as ge
g = ge.load_package()
dims, B = [4096, 4096, 4096, 1000], 4096
net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                     flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
rng = np.random.default_rng(1)
x = rng.random((B, dims[0]), dtype=np.float32)
lab = rng.integers(0, 1000, B).astype(np.int32)
for it in range(3):
    print("--- step", it, file=sys.stderr, flush=True)
    net.gradientStepLabeled(x, lab, 0.01, 0.9)
net.sync()
