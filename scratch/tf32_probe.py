"""TF32 bring-up probe: bare contraction and per-layer update norms of one step (run on the GPU box)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
g = ge.load_package()
tag = "TMAP=%s DBG=%s" % (os.environ.get("B2NN_TF32_TMAP"), os.environ.get("B2NN_GEMM_DBG"))
M, K, N = 128, 64, 256
net = g.GeneralNeuralNet([K, N], g.IDENTITY, g.IDENTITY_PRIME, g.IDENTITY, g.IDENTITY_PRIME, g.HALF_SQUARED_ERROR,
                         g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_TF32)
rng = np.random.RandomState(0)
W = rng.randn(K, N) * 0.5
x = rng.randn(M, K)
net.setWeights(W.ravel())
y = net.propagate(x)
ref = x @ W
print(tag, "gemm max|y| %.4g max|ref| %.4g err %.4g  y[0,:4]=%s ref[0,:4]=%s nan=%d" %
      (np.abs(y).max(), np.abs(ref).max(), np.abs(y - ref).max(), y[0, :4], ref[0, :4], np.isnan(y).sum()))
# ones: every product is exactly representable
net.setWeights(np.ones(K * N))
y1 = net.propagate(np.ones((M, K)))
print(tag, "ones: min %.4g max %.4g (expect %d)" % (y1.min(), y1.max(), K))
dims = [256, 128, 64, 32]
n3 = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_TF32,
                                    flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
w0 = n3.getWeights()
xb = rng.rand(64, 256).astype(np.float32)
lab = rng.randint(0, 32, 64).astype(np.int32)
loss = n3.gradientStepLabeled(xb, lab, 0.05, 0.9, returnLoss=True)
du = n3.getWeights() - w0
off = 0
for l in range(3):
    n = dims[l] * dims[l + 1]
    print(tag, "layer %d |dW| %.4g" % (l, np.linalg.norm(du[off:off + n])))
    off += n
print(tag, "loss %.5f (ln 32 = %.5f)" % (loss, np.log(32)))
