"""does a single-GPU TF32 net update for every batch size?  (tests/test_dp.py tf32-nvls-8: the single-GPU comparator at batch
2048 came back with unchanged weights)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
g = ge.load_package()
dims = [256, 384, 128, 40]
flags = g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX
rng = np.random.RandomState(11)
for prec, name in ((g.capi.PREC_TF32, "tf32"), (g.capi.PREC_BF16, "bf16")):
    for B in (256, 512, 1024, 2048, 4096):
        x = rng.rand(B, dims[0]).astype(np.float32)
        lab = rng.randint(0, dims[-1], B).astype(np.int32)
        net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=prec, flags=flags)
        w0 = net.getWeights()
        losses = [net.gradientStepLabeled(x, lab, 0.05, 0.9, returnLoss=True) for _ in range(3)]
        w = net.getWeights()
        off, per = 0, []
        for l in range(3):
            n = dims[l] * dims[l + 1]
            per.append(float(np.abs(w[off:off + n] - w0[off:off + n]).max()))
            off += n
        print(name, "B", B, "losses", ["%.4f" % l for l in losses], "max|dW| per layer", ["%.3g" % p for p in per], flush=True)
        net.close()
