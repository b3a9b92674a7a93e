"""GPU tests of the TF32 tcgen05 path (B2NN_PREC_TF32: FP32 activations / deltas / master weights read by the
tensor cores as TF32, FP32 accumulation) through the C-ABI, against the FP64 CPU oracle.

Tolerances (stated here): TF32 operands carry a 10-bit mantissa --
  - plain contraction y = x W (identity activations): |diff| <= 5e-5 * max|y| against operands rounded to nearest
    TF32 (TMA's TFLOAT32 map type rounds the FP32 words on their way into shared memory; measured <= 6e-6, what is
    left is the FP32 accumulation order);
  - one-step update: <= 1e-3 relative (Frobenius, per layer; measured <= 1.3e-4) against a numpy model that rounds
    to TF32 where the kernels do; <= 3e-2 and cosine >= 0.999 against the FP64 oracle (measured <= 1.9e-2 in the
    layers below a leaky-ReLU: a unit whose pre-activation lies inside the TF32 noise takes the other slope;
    <= 6e-4 in the top layer and with smooth activations); batch loss within 1e-4 relative;
  - loss curve over the first 40 steps within 1 % of the oracle's (measured 0.3 %).
Measured values: profiles/r01_tf32_parity_report.txt.  The file sorts after the BF16 tests on purpose: newest path.
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REPORT = os.environ.get("B2NN_TEST_REPORT")


def report(line):
    """measured errors, kept when B2NN_TEST_REPORT names a file (used to tighten the bounds above)"""
    if REPORT:
        with open(REPORT, "a") as f:
            f.write(line + "\n")


def tf32_round(a):
    """round-to-nearest-even to TF32 (8-bit exponent, 10-bit mantissa), returned as float64"""
    a32 = np.ascontiguousarray(a, dtype=np.float32)
    u = a32.view(np.uint32).astype(np.uint64)
    u = (u + 0xFFF + ((u >> 13) & 1)) & 0xFFFFE000
    return u.astype(np.uint32).view(np.float32).astype(np.float64)


def tf32_net(g, dims, **kw):
    flags = g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX
    return g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_TF32,
                                          flags=flags, **kw)


@pytest.mark.parametrize("M,K,N", [(128, 32, 256), (256, 512, 512), (100, 200, 1000), (4096, 1000, 136),
                                    (37, 8, 8), (300, 2048, 264), (2560, 96, 2048)])
def test_tf32_gemm_forward_identity(g, M, K, N):
    """GeneralNeuralNet with identity activations makes propagate() the bare kind::tf32 contraction y = x W
    (A K-major, B MN-major 32 x 32 boxes): ragged M/N/K, zero fill, descriptors, TMEM epilogue."""
    net = g.GeneralNeuralNet([K, N], g.IDENTITY, g.IDENTITY_PRIME, g.IDENTITY, g.IDENTITY_PRIME,
                             g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_TF32)
    rng = np.random.RandomState(M + K + N)
    W = rng.randn(K, N) * 0.5
    net.setWeights(W.ravel())
    x = rng.randn(M, K)
    y = net.propagate(x)
    ref = tf32_round(x) @ tf32_round(W)          # rounded operands, FP64 accumulate
    err = np.abs(y - ref).max()
    report("gemm %dx%dx%d err/max %.3g" % (M, K, N, err / max(1.0, np.abs(ref).max())))
    assert err <= 5e-5 * max(1.0, np.abs(ref).max()), (err, np.abs(ref).max())


def test_tf32_weights_roundtrip(g):
    net = tf32_net(g, [64, 40, 24])
    v = np.linspace(-1, 1, net.numWeights)
    net.setWeights(v)
    assert np.array_equal(net.getWeights(), v.astype(np.float32).astype(np.float64))   # FP32 master, untouched
    assert np.array_equal(net.getVelocity(), np.zeros(net.numWeights))


def emulate_tf32_step(dims, w0, x, lab, step, alpha=0.01, act="leaky"):
    """numpy model of the TF32 path: FP32 arrays in memory, every GEMM operand rounded to TF32 on its way into the
    tensor core, products accumulated in FP64 (the kernels accumulate in FP32).  Returns (mean loss, update
    v = step*G/B per layer)."""
    L = len(dims) - 1
    B = x.shape[0]
    f32 = lambda a: np.asarray(a, dtype=np.float32).astype(np.float64)
    W, off = [], 0
    for l in range(L):
        n = dims[l] * dims[l + 1]
        W.append(w0[off:off + n].reshape(dims[l], dims[l + 1]))
        off += n
    if act == "sigmoid":   # smooth: no slope flips (f' from the stored activation)
        leaky, fprime = (lambda z: 1.0 / (1.0 + np.exp(-z))), (lambda a: a * (1.0 - a))
    else:
        leaky, fprime = (lambda z: np.where(z > 0, z, alpha * z)), (lambda a: np.where(a > 0, 1.0, alpha))
    a = [f32(leaky(f32(x)))]
    for l in range(L - 1):
        a.append(f32(leaky(tf32_round(a[l]) @ tf32_round(W[l]))))
    z = f32(tf32_round(a[L - 1]) @ tf32_round(W[L - 1]))
    zs = z - z.max(axis=1, keepdims=True)
    lse = np.log(np.exp(zs).sum(axis=1, keepdims=True))
    p = np.exp(zs - lse)
    loss = float(np.mean(-(zs - lse)[np.arange(B), lab]))
    t = np.zeros_like(p)
    t[np.arange(B), lab] = 1.0
    d = [None] * (L + 1)
    d[L] = f32(p - t)
    for l in range(L - 1, 0, -1):
        d[l] = f32((tf32_round(d[l + 1]) @ tf32_round(W[l]).T) * fprime(a[l]))
    return loss, [step * (tf32_round(a[l]).T @ tf32_round(d[l + 1])) / B for l in range(L)]


@pytest.mark.parametrize("dims,B", [([512, 384, 256, 100], 256), ([784, 300, 100, 10], 64),
                                     ([1000, 520, 1000], 136)])
def test_tf32_one_step_vs_model_and_oracle(g, orc, dims, B):
    """forward (K-major x MN-major), delta (K-major x K-major) and gradient (MN-major x MN-major, K = batch)
    contractions plus the FP32 epilogues and the momentum update, one step"""
    net = tf32_net(g, dims)
    ref = orc.OracleNet(dims)
    ref.set_weights(net.getWeights())
    rng = np.random.RandomState(1)
    x = rng.rand(B, dims[0])
    lab = rng.randint(0, dims[-1], B).astype(np.int32)
    t = np.zeros((B, dims[-1]))
    t[np.arange(B), lab] = 1.0
    w0 = ref.get_weights()
    ref_loss = np.mean([ref.loss(x[i], t[i]) for i in range(B)])
    emu_loss, emu_v = emulate_tf32_step(dims, w0, x, lab, 0.05)
    loss = net.gradientStepLabeled(x.astype(np.float32), lab, 0.05, 0.9, returnLoss=True)
    ref.gradient_step(x, t, 0.05, 0.9)
    report("step %s loss %.6f oracle %.6f model %.6f" % (dims, loss, ref_loss, emu_loss))
    assert abs(loss - ref_loss) <= 1e-4 * max(1.0, abs(ref_loss))
    assert abs(loss - emu_loss) <= 1e-4 * max(1.0, abs(emu_loss))
    du, du_ref = w0 - net.getWeights(), w0 - ref.get_weights()
    v = net.getVelocity()
    off = 0
    for l in range(len(dims) - 1):
        n = dims[l] * dims[l + 1]
        a, r, e = du[off:off + n], du_ref[off:off + n], emu_v[l].ravel()
        rel_e, rel_r = np.linalg.norm(a - e) / np.linalg.norm(e), np.linalg.norm(a - r) / np.linalg.norm(r)
        report("step %s layer %d: vs model %.3g, vs oracle %.3g" % (dims, l, rel_e, rel_r))
        assert rel_e <= 1e-3, (l, rel_e)
        assert rel_r <= 3e-2, (l, rel_r)
        assert a @ r >= 0.999 * np.linalg.norm(a) * np.linalg.norm(r)
        assert np.linalg.norm(v[off:off + n] - e) <= 1e-3 * np.linalg.norm(e)
        off += n


def test_tf32_general_net_sigmoid_vs_oracle(g, orc):
    """GeneralNeuralNet semantics (G:183-221, G:251-307, G:317-366): sigmoid hidden layers, input activation (Q2),
    sigmoid output, half squared error with dense targets; three momentum steps within 2e-3 of the FP64 oracle."""
    dims, B = [320, 256, 192, 24], 128
    net = g.GeneralNeuralNet(dims, g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME, g.HALF_SQUARED_ERROR,
                             g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_TF32, flags=g.capi.FLAG_SCALED_INIT)
    ref = orc.OracleNet(dims, kind=orc.NET_GENERAL, inner_act="sigmoid", inner_param=0.0, last_act="sigmoid")
    ref.set_weights(net.getWeights())
    rng = np.random.RandomState(3)
    x = tf32_round(rng.rand(B, dims[0]) - 0.3)
    t = rng.rand(B, dims[-1])
    w0 = ref.get_weights()
    for it in range(3):
        ref_loss = sum(ref.loss(x[i], t[i]) for i in range(B)) / B
        loss = net.gradientStep((x, t), 0.5, 0.9, returnLoss=True)
        ref.gradient_step(x, t, 0.5, 0.9)
        assert abs(loss - ref_loss) <= 2e-3 * max(1.0, abs(ref_loss)), (it, loss, ref_loss)
    du, du_ref = w0 - net.getWeights(), w0 - ref.get_weights()
    off = 0
    for l in range(len(dims) - 1):
        n = dims[l] * dims[l + 1]
        a, r = du[off:off + n], du_ref[off:off + n]
        rel = np.linalg.norm(a - r) / np.linalg.norm(r)
        report("general sigmoid layer %d: vs oracle %.3g" % (l, rel))
        assert rel <= 2e-3, (l, rel)
        off += n


def test_tf32_loss_curve_and_eval(g, orc):
    dims, B, n, iters = [256, 192, 128, 16], 128, 2048, 40
    net = tf32_net(g, dims)
    ref = orc.OracleNet(dims)
    ref.set_weights(net.getWeights())
    rng = np.random.RandomState(3)
    centers = rng.rand(16, 256)
    lab = rng.randint(0, 16, n).astype(np.int32)
    x = np.clip(centers[lab] + 0.3 * rng.randn(n, 256), 0, 1)
    t = np.zeros((n, 16))
    t[np.arange(n), lab] = 1.0
    ds = g.Dataset(net, x.astype(np.float32), labels=lab)
    idx = g.Sampler(n, seed=1).draw(iters, B)
    losses = np.empty(iters)
    g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), iters, B,
                                                 0.05, 0.9, losses.ctypes.data_as(C.POINTER(C.c_double))))
    ref_losses = ref.train(x, t, iters, B, 0.05, 0.9, idx=idx, want_loss=True)
    report("loss curve max rel %.3g" % np.max(np.abs(losses - ref_losses) / np.maximum(1.0, np.abs(ref_losses))))
    assert np.all(np.abs(losses - ref_losses) <= 1e-2 * np.maximum(1.0, np.abs(ref_losses)))
    assert losses[-1] < losses[0]
    hits, mean_loss, pred = ds.evaluate(want_pred=True)
    rhits, rpred = ref.evaluate(x, lab)
    assert abs(hits - rhits) <= 0.02 * n
    assert (pred == rpred).mean() >= 0.97
    assert abs(mean_loss - ref.validate(x, t, n)) <= 2e-2 * max(1.0, mean_loss)


def test_tf32_pixel_byte_rows_match_float_rows(g):
    """rows kept as IDX pixel bytes in pinned host memory (B2NN_DT_U8, (raw & 0xff) / 255.0 as in M:97) against the
    same rows converted on the host like the reference's loader and staged on the device as FP32"""
    dims, B = [256, 128, 32], 64
    a, b = tf32_net(g, dims), tf32_net(g, dims)
    rng = np.random.RandomState(6)
    raw = rng.randint(0, 256, (512, 256)).astype(np.uint8)
    x = (raw.astype(np.float64) / 255.0).astype(np.float32)
    lab = rng.randint(0, 32, 512).astype(np.int32)
    da = g.Dataset(a, x, labels=lab, location=g.capi.DATA_DEVICE)
    db = g.Dataset(b, raw, labels=lab, location=g.capi.DATA_HOST_PINNED)
    idx = g.Sampler(512, seed=1).draw(6, B)
    for net, ds in ((a, da), (b, db)):
        g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), 6, B,
                                                     0.05, 0.9, None))
    assert np.array_equal(a.getWeights(), b.getWeights())
    assert da.evaluate()[0] == db.evaluate()[0]


def test_tf32_peer_memory_export_without_communicator(g):
    """TF32 nets take the fused peer-memory / multicast exchange too (the kernel replicates their FP32 weights instead
    of a BF16 copy; 2-GPU parity in tests/test_dp.py): the CUDA IPC handle of the arena exports, the multicast export
    still needs a communicator (world > 1) and says so without touching the net"""
    net = tf32_net(g, [64, 32, 8])
    assert len(net.commIpcExport()) == 64
    assert net.commNvlsExport() is None
    w = net.getWeights()
    assert np.isfinite(w).all() and w.size == net.numWeights
