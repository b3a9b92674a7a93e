"""GPU tests of the BF16 tcgen05 path through the C-ABI, against the FP64 CPU oracle.

Tolerances (stated here): BF16 operands (8-bit mantissa) with FP32 accumulation --
  - plain contraction y = x W (identity activations): |diff| <= 2e-2 * max|y|;
  - one-step update: <= 1e-2 relative (Frobenius, per layer) against a numpy model that rounds to
    BF16 where the kernels do; <= 1e-1 and cosine >= 0.995 against the FP64 oracle (leaky-ReLU slope
    flips near zero dominate, see test_one_step_vs_bf16_model_and_oracle); batch loss within 2e-2;
  - loss curve over the first 40 steps within 5 % of the oracle's.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bf16_round(a):
    """round-to-nearest-even to bfloat16, returned as float64"""
    a32 = np.ascontiguousarray(a, dtype=np.float32)
    u = a32.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).astype(np.float64)


def wide_net(g, dims, **kw):
    flags = g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX
    return g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                          flags=flags, **kw)


@pytest.mark.parametrize("M,K,N", [(128, 64, 256), (256, 512, 512), (100, 200, 1000), (4096, 1000, 136),
                                    (37, 8, 8), (300, 4096, 264), (2560, 128, 2048)])
def test_gemm_forward_identity(g, M, K, N):
    """GeneralNeuralNet with identity activations makes propagate() the bare tcgen05 contraction
    y = x W: exercises TMA zero-fill on ragged M/N/K, the UMMA descriptors and the TMEM epilogue."""
    net = g.GeneralNeuralNet([K, N], g.IDENTITY, g.IDENTITY_PRIME, g.IDENTITY, g.IDENTITY_PRIME,
                             g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_BF16)
    rng = np.random.RandomState(M + K + N)
    W = rng.randn(K, N) * 0.5
    net.setWeights(W.ravel())
    x = rng.randn(M, K)
    y = net.propagate(x)
    ref = bf16_round(x) @ bf16_round(W)          # exact operands, FP64 accumulate
    err = np.abs(y - ref).max()
    assert err <= 1e-3 * max(1.0, np.abs(ref).max()), (err, np.abs(ref).max())


def test_gemm_two_layer_chain(g):
    """two chained contractions: the first epilogue stores BF16 activations that feed the second"""
    K, H, N, M = 320, 520, 72, 260
    net = g.GeneralNeuralNet([K, H, N], g.IDENTITY, g.IDENTITY_PRIME, g.IDENTITY, g.IDENTITY_PRIME,
                             g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_BF16)
    rng = np.random.RandomState(9)
    W0, W1 = rng.randn(K, H) * 0.1, rng.randn(H, N) * 0.1
    net.setWeights(np.concatenate([W0.ravel(), W1.ravel()]))
    x = rng.randn(M, K)
    h = bf16_round(bf16_round(x) @ bf16_round(W0))
    ref = h @ bf16_round(W1)
    y = net.propagate(x)
    assert np.abs(y - ref).max() <= 2e-2 * np.abs(ref).max()


def test_weights_roundtrip_and_init(g, orc):
    dims = [64, 40, 24]
    net = wide_net(g, dims)
    ref = orc.OracleNet(dims)
    w = net.getWeights()
    r = ref.get_weights()
    # scaled init: (U[-0.5,0.5) from the same LCG stream) * 2*sqrt(6/fan_in), stored as FP32
    sc = np.concatenate([np.full(64 * 40, 2 * np.sqrt(6 / 64)), np.full(40 * 24, 2 * np.sqrt(6 / 40))])
    assert np.allclose(w, (r * sc).astype(np.float32), rtol=0, atol=0)
    v = np.linspace(-1, 1, net.numWeights)
    net.setWeights(v)
    assert np.array_equal(net.getWeights(), v.astype(np.float32).astype(np.float64))


def emulate_bf16_step(dims, w0, x, lab, step, alpha=0.01, act="leaky"):
    """numpy model of the BF16 path: operands rounded to bfloat16 exactly where the kernels round
    them (a_l, delta_l, the two BF16 weight copies), products accumulated in FP64 (the kernels
    accumulate in FP32), FP32 master update.  Returns (mean loss, update v = step*G/B per layer)."""
    L = len(dims) - 1
    B = x.shape[0]
    W, off = [], 0
    for l in range(L):
        n = dims[l] * dims[l + 1]
        W.append(w0[off:off + n].reshape(dims[l], dims[l + 1]))
        off += n
    if act == "sigmoid":   # smooth: no slope flips (f' from the stored activation)
        leaky, fprime = (lambda z: 1.0 / (1.0 + np.exp(-z))), (lambda a: a * (1.0 - a))
    else:
        leaky, fprime = (lambda z: np.where(z > 0, z, alpha * z)), (lambda a: np.where(a > 0, 1.0, alpha))
    a = [bf16_round(leaky(x.astype(np.float32).astype(np.float64)))]
    for l in range(L - 1):
        a.append(bf16_round(leaky(a[l] @ bf16_round(W[l]))))
    z = (a[L - 1] @ bf16_round(W[L - 1])).astype(np.float32).astype(np.float64)
    zs = z - z.max(axis=1, keepdims=True)
    lse = np.log(np.exp(zs).sum(axis=1, keepdims=True))
    p = np.exp(zs - lse)
    loss = float(np.mean(-(zs - lse)[np.arange(B), lab]))
    t = np.zeros_like(p)
    t[np.arange(B), lab] = 1.0
    d = [None] * (L + 1)
    d[L] = bf16_round(p - t)
    for l in range(L - 1, 0, -1):
        d[l] = bf16_round((d[l + 1] @ bf16_round(W[l]).T) * fprime(a[l]))
    return loss, [step * (a[l].T @ d[l + 1]) / B for l in range(L)]


@pytest.mark.parametrize("dims,B", [([512, 384, 256, 100], 256), ([784, 300, 100, 10], 64),
                                     ([1000, 520, 1000], 136)])
def test_one_step_vs_bf16_model_and_oracle(g, orc, dims, B):
    """Tight check against a numpy model that rounds to BF16 at the same points (<= 1e-2 per layer),
    loose check against the FP64 oracle: with leaky-ReLU a unit whose pre-activation is within the
    BF16 noise of zero takes the other slope, so a fraction f of delta entries is off by ~100 % and
    the Frobenius error is ~sqrt(f) (a few %); hence <= 1e-1 per layer plus cosine >= 0.995."""
    net = wide_net(g, dims)
    ref = orc.OracleNet(dims)
    ref.set_weights(net.getWeights())
    rng = np.random.RandomState(1)
    x = rng.rand(B, dims[0])
    lab = rng.randint(0, dims[-1], B).astype(np.int32)
    t = np.zeros((B, dims[-1]))
    t[np.arange(B), lab] = 1.0
    w0 = ref.get_weights()
    ref_loss = np.mean([ref.loss(x[i], t[i]) for i in range(B)])
    emu_loss, emu_v = emulate_bf16_step(dims, w0, x, lab, 0.05)
    loss = net.gradientStepLabeled(x.astype(np.float32), lab, 0.05, 0.9, returnLoss=True)
    ref.gradient_step(x, t, 0.05, 0.9)
    assert abs(loss - ref_loss) <= 2e-2 * max(1.0, abs(ref_loss))
    assert abs(loss - emu_loss) <= 1e-3 * max(1.0, abs(emu_loss))
    du, du_ref = w0 - net.getWeights(), w0 - ref.get_weights()
    v = net.getVelocity()
    off = 0
    for l in range(len(dims) - 1):
        n = dims[l] * dims[l + 1]
        a, r, e = du[off:off + n], du_ref[off:off + n], emu_v[l].ravel()
        assert np.linalg.norm(a - e) <= 1e-2 * np.linalg.norm(e), (l, np.linalg.norm(a - e) / np.linalg.norm(e))
        assert np.linalg.norm(a - r) <= 1e-1 * np.linalg.norm(r), (l, np.linalg.norm(a - r) / np.linalg.norm(r))
        assert a @ r >= 0.995 * np.linalg.norm(a) * np.linalg.norm(r)
        assert np.linalg.norm(v[off:off + n] - e) <= 1e-2 * np.linalg.norm(e)
        off += n


@pytest.mark.parametrize("inner", ["sigmoid", "tanh", "relu"])
def test_general_net_smooth_activations_vs_oracle(g, orc, inner):
    """GeneralNeuralNet semantics (G:183-221, G:251-307, G:317-366) on the tensor-core path: `inner` activation in the
    hidden layers AND on the input (Q2), sigmoid output, half squared error with dense targets.  Smooth activations
    have no slope flips, so three momentum steps stay within 3e-2 (Frobenius, per layer) of the FP64 oracle and the
    batch loss within 2e-2; ReLU flips like leaky-ReLU: 1e-1 and cosine >= 0.99."""
    dims, B = [320, 256, 192, 24], 128
    acts = {"sigmoid": (g.SIGMOID, g.SIGMOID_PRIME), "tanh": (g.TANH, g.TANH_PRIME), "relu": (g.RELU, g.RELU_PRIME)}
    f, fp = acts[inner]
    net = g.GeneralNeuralNet(dims, f, fp, g.SIGMOID, g.SIGMOID_PRIME, g.HALF_SQUARED_ERROR,
                             g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_BF16, flags=g.capi.FLAG_SCALED_INIT)
    ref = orc.OracleNet(dims, kind=orc.NET_GENERAL, inner_act=inner, inner_param=0.0, last_act="sigmoid")
    ref.set_weights(net.getWeights())
    rng = np.random.RandomState(3)
    x = bf16_round(rng.rand(B, dims[0]) - 0.3)          # exactly representable inputs of both signs
    t = rng.rand(B, dims[-1])
    w0 = ref.get_weights()
    for it in range(3):
        ref_loss = sum(ref.loss(x[i], t[i]) for i in range(B)) / B
        loss = net.gradientStep((x, t), 0.5, 0.9, returnLoss=True)
        ref.gradient_step(x, t, 0.5, 0.9)
        assert abs(loss - ref_loss) <= 2e-2 * max(1.0, abs(ref_loss)), (it, loss, ref_loss)
    du, du_ref = w0 - net.getWeights(), w0 - ref.get_weights()
    tol, cos_min = (1e-1, 0.99) if inner == "relu" else (3e-2, 0.999)
    off = 0
    for l in range(len(dims) - 1):
        n = dims[l] * dims[l + 1]
        a, r = du[off:off + n], du_ref[off:off + n]
        assert np.linalg.norm(a - r) <= tol * np.linalg.norm(r), (l, np.linalg.norm(a - r) / np.linalg.norm(r))
        assert a @ r >= cos_min * np.linalg.norm(a) * np.linalg.norm(r)
        off += n


def test_dense_target_step_matches_labeled(g):
    dims, B = [256, 128, 32], 64
    a, b = wide_net(g, dims), wide_net(g, dims)
    rng = np.random.RandomState(2)
    x = rng.rand(B, 256)
    lab = rng.randint(0, 32, B).astype(np.int32)
    t = np.zeros((B, 32))
    t[np.arange(B), lab] = 1.0
    la = a.gradientStep((x, t), 0.05, 0.9, returnLoss=True)       # drop-in double path
    lb = b.gradientStepLabeled(x, lab, 0.05, 0.9, returnLoss=True)
    assert abs(la - lb) <= 1e-6 * max(1.0, abs(lb))
    assert np.array_equal(a.getWeights(), b.getWeights())


def test_host_pinned_bf16_rows_match_device(g):
    """rows staged on the host as BF16 bit patterns (what bench.py's e2e leg feeds) against the same values resident
    on the device as FP32; inputs of both signs so that the input activation f(x) (S:183-186) is exercised"""
    dims, B = [256, 128, 32], 64
    a, b = wide_net(g, dims), wide_net(g, dims)
    rng = np.random.RandomState(5)
    x = (rng.rand(512, 256).astype(np.float32) - 0.5)
    u = x.view(np.uint32)
    bits = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)          # round to nearest even
    x_rounded = (bits.astype(np.uint32) << 16).view(np.float32)
    lab = rng.randint(0, 32, 512).astype(np.int32)
    da = g.Dataset(a, x_rounded, labels=lab, location=g.capi.DATA_DEVICE)
    db = g.Dataset(b, bits, labels=lab, location=g.capi.DATA_HOST_PINNED)
    idx = g.Sampler(512, seed=1).draw(6, B)
    import ctypes as C
    for net, ds in ((a, da), (b, db)):
        g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), 6, B,
                                                     0.05, 0.9, None))
    assert np.array_equal(a.getWeights(), b.getWeights())


def test_loss_curve_and_eval(g, orc):
    dims, B, n, iters = [256, 192, 128, 16], 128, 2048, 40
    net = wide_net(g, dims)
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
    import ctypes as C
    losses = np.empty(iters)
    g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), iters, B,
                                                 0.05, 0.9, losses.ctypes.data_as(C.POINTER(C.c_double))))
    ref_losses = ref.train(x, t, iters, B, 0.05, 0.9, idx=idx, want_loss=True)
    assert np.all(np.abs(losses - ref_losses) <= 5e-2 * np.maximum(1.0, np.abs(ref_losses)))
    assert losses[-1] < losses[0]
    hits, mean_loss, pred = ds.evaluate(want_pred=True)
    rhits, rpred = ref.evaluate(x, lab)
    assert abs(hits - rhits) <= 0.03 * n                    # two independently trained copies in different precision
    assert (pred == rpred).mean() >= 0.95
    assert abs(mean_loss - ref.validate(x, t, n)) <= 5e-2 * max(1.0, mean_loss)


def test_eval_argmax_last_max_wins(g):
    dims = [64, 16]
    net = wide_net(g, dims)
    net.setWeights(np.zeros(net.numWeights))
    x = np.random.RandomState(0).rand(40, 64).astype(np.float32)
    ds = g.Dataset(net, x, labels=np.full(40, 15, dtype=np.int32))
    hits, _, pred = ds.evaluate(want_pred=True)
    assert hits == 40 and (pred == 15).all()                # MNISTTrainer.java:166-168


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_eval_argmax_compares_activated_outputs(g, prec):
    """M:164-168 compares the ACTIVATED outputs with `>=` from index 0: with a ReLU output layer every row whose
    logits are all negative ties at 0 and the LAST index wins (argmax of the logits would pick another class); rows
    with a positive logit keep its index; equal positive outputs -> the later index."""
    dL, d0 = 16, 64
    net = g.GeneralNeuralNet([d0, dL], g.IDENTITY, g.IDENTITY_PRIME, g.RELU, g.RELU_PRIME, g.HALF_SQUARED_ERROR,
                             g.HALF_SQUARED_ERROR_PRIME,
                             precision=g.capi.PREC_BF16 if prec == "bf16" else g.capi.PREC_TF32)
    W = np.zeros((d0, dL))
    W[0, :] = -np.arange(1, dL + 1) / 8.0        # logits of rows with x[0] = 1: all negative, largest at index 0
    W[1, 3] = 2.0                                # rows with x[1] = 1 as well: index 3 positive
    W[2, 5] = W[2, 9] = 1.0                      # rows with x[2] = 1: indices 5 and 9 tie at a positive value
    net.setWeights(W.ravel())
    x = np.zeros((48, d0), dtype=np.float32)
    x[:, 0] = 1.0
    x[16:32, 1] = 1.0
    x[32:, 0] = 0.0
    x[32:, 2] = 4.0
    lab = np.concatenate([np.full(16, dL - 1), np.full(16, 3), np.full(16, 9)]).astype(np.int32)
    ds = g.Dataset(net, x, labels=lab)
    hits, _, pred = ds.evaluate(want_pred=True)
    assert (pred == lab).all() and hits == 48, pred
    y = net.propagate(x.astype(np.float64))
    ref = np.array([max(i for i in range(dL) if r[i] >= r.max()) for r in y])
    assert (pred == ref).all()


@pytest.mark.parametrize("kind", ["softmax", "general"])
def test_propagate_output_pass_on_device(g, kind):
    """NeuralNet.propagate (src/NeuralNet.java:16) of a tensor-core net: softmax (S:357-376) / last activation
    (G:214-218) of the FP32 logits is computed on the device in double -- equal to numpy on the same logits"""
    d0, dL, n = 96, 40, 50
    rng = np.random.RandomState(12)
    W = bf16_round(rng.randn(d0, dL) * 0.3)
    x = bf16_round(rng.rand(n, d0))
    if kind == "softmax":
        net = g.SoftmaxCrossEntropyNeuralNet([d0, dL], g.IDENTITY, g.IDENTITY_PRIME, precision=g.capi.PREC_BF16,
                                             flags=g.capi.FLAG_STABLE_SOFTMAX)
    else:
        net = g.GeneralNeuralNet([d0, dL], g.IDENTITY, g.IDENTITY_PRIME, g.TANH, g.TANH_PRIME, g.HALF_SQUARED_ERROR,
                                 g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_BF16)
    net.setWeights(W.ravel())
    z = (x @ W).astype(np.float32).astype(np.float64)
    if kind == "softmax":
        e = np.exp(z - z.max(axis=1, keepdims=True))
        ref = e / e.sum(axis=1, keepdims=True)
    else:
        ref = np.tanh(z)
    y = net.propagate(x)
    assert np.abs(y - ref).max() <= 5e-6    # FP32 accumulation order of the logits; the output pass itself is double
    if kind == "softmax":
        assert np.abs(y.sum(axis=1) - 1.0).max() <= 1e-12


@pytest.mark.parametrize("d0", [256, 100])   # 100: rows not 8-byte aligned, scalar load path
def test_pixel_byte_rows_match_float_rows(g, d0):
    """rows kept as IDX pixel bytes in pinned host memory (B2NN_DT_U8, value = (raw & 0xff) / 255.0 as in M:97) against
    the same rows converted on the host the way the reference's loader does and staged on the device as FP32"""
    dims, B = [d0, 128, 32], 64
    a, b = wide_net(g, dims), wide_net(g, dims)
    rng = np.random.RandomState(6)
    raw = rng.randint(0, 256, (512, d0)).astype(np.uint8)
    x = (raw.astype(np.float64) / 255.0).astype(np.float32)
    lab = rng.randint(0, 32, 512).astype(np.int32)
    da = g.Dataset(a, x, labels=lab, location=g.capi.DATA_DEVICE)
    db = g.Dataset(b, raw, labels=lab, location=g.capi.DATA_HOST_PINNED)
    idx = g.Sampler(512, seed=1).draw(6, B)
    import ctypes as C
    for net, ds in ((a, da), (b, db)):
        g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), 6, B,
                                                     0.05, 0.9, None))
    assert np.array_equal(a.getWeights(), b.getWeights())
    assert da.evaluate()[0] == db.evaluate()[0]


def test_host_pinned_dataset_matches_device(g):
    dims, B = [256, 128, 32], 64
    a, b = wide_net(g, dims), wide_net(g, dims)
    rng = np.random.RandomState(4)
    x = rng.rand(512, 256).astype(np.float32)
    lab = rng.randint(0, 32, 512).astype(np.int32)
    da = g.Dataset(a, x, labels=lab, location=g.capi.DATA_DEVICE)
    db = g.Dataset(b, x, labels=lab, location=g.capi.DATA_HOST_PINNED)
    idx = g.Sampler(512, seed=1).draw(6, B)
    import ctypes as C
    for net, ds in ((a, da), (b, db)):
        g.capi.check(g.capi.lib().b2nn_train_indexed(net._h, ds._h, idx.ctypes.data_as(C.POINTER(C.c_int32)), 6, B,
                                                     0.05, 0.9, None))
    assert np.array_equal(a.getWeights(), b.getWeights())


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
@pytest.mark.parametrize("dims,B", [([512, 384, 256, 100], 256), ([300, 131, 70, 10], 96)])
def test_fused_update_epilogue_matches_separate_update_kernel(g, prec, dims, B, monkeypatch):
    """single GPU: the momentum update rides in the gradient GEMM's epilogue (EPI_DW_UPDATE: accumulator chunk staged
    through shared memory, coalesced read-modify-write of w / v, S:325-342) by default; B2NN_FUSED_UPDATE=0 stores the
    summed gradient and runs the separate update kernel.  Same FP32 operations on the same accumulators: weights and
    velocity agree to FP32 rounding of one multiply-add (<= 1e-6 of the update, the compiler may contract differently)
    over 4 momentum steps; dims with widths that are no multiple of 4 exercise the ragged right edge of a quad."""
    precision = g.capi.PREC_BF16 if prec == "bf16" else g.capi.PREC_TF32
    flags = g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX
    rng = np.random.RandomState(17)
    x = rng.rand(B, dims[0]).astype(np.float32)
    lab = rng.randint(0, dims[-1], B).astype(np.int32)
    out = []
    for fused in ("1", "0"):
        monkeypatch.setenv("B2NN_FUSED_UPDATE", fused)
        net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=precision, flags=flags)
        w0 = net.getWeights()
        k0 = net.kernelLaunches()
        for _ in range(4):
            net.gradientStepLabeled(x, lab, 0.05, 0.9)
        out.append((net.getWeights() - w0, net.getVelocity(), net.kernelLaunches() - k0))
        net.close()
    (du_f, v_f, k_f), (du_s, v_s, k_s) = out
    assert k_f == k_s - 4 * (len(dims) - 1)                 # one update kernel per layer and step fewer
    assert np.linalg.norm(du_f - du_s) <= 1e-6 * np.linalg.norm(du_s)
    assert np.linalg.norm(v_f - v_s) <= 1e-6 * np.linalg.norm(v_s)


def test_profile_windows_per_launch_and_per_run(g):
    """bench.py's roofline instrument (b2nn_profile_enable): mode 1 brackets every contraction launch with its own event
    pair, mode 2 brackets each RUN of consecutive contraction launches (the forward GEMMs; the backward GEMMs).  Both
    count the same launches and the same algorithmic FLOPs (2 M N K per launch), neither changes what the step
    computes, and the per-run windows are fewer than the launches."""
    dims, B, steps = [512, 384, 256, 100], 256, 3
    L = len(dims) - 1
    rng = np.random.RandomState(23)
    x = rng.rand(B, dims[0]).astype(np.float32)
    lab = rng.randint(0, dims[-1], B).astype(np.int32)
    flops_step = 2.0 * B * (3 * sum(a * b for a, b in zip(dims[:-1], dims[1:])) - dims[0] * dims[1])
    res = {}
    for mode in (0, 1, 2):
        net = wide_net(g, dims)
        if mode:
            net.profileEnable(mode)
        for _ in range(steps):
            net.gradientStepLabeled(x, lab, 0.05, 0.9)
        if mode:
            ms, n, fl = net.profileRead()
            cls = net.profileReadClasses()
            net.profileEnable(False)
            assert n == steps * (3 * L - 1) and fl == steps * flops_step
            assert ms > 0 and abs(cls["contraction"][0] - ms) <= 1e-9 * ms
            assert cls["output"][1] == steps
            res[mode] = (net.getWeights(), cls["contraction"][1])
        else:
            res[mode] = (net.getWeights(), 0)
        net.close()
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][0], res[2][0])
    assert res[1][1] == steps * (3 * L - 1)      # a window per launch
    assert res[2][1] == 2 * steps                # forward run + backward run per step


def test_profile_in_kernel_stamps(g):
    """b2nn_profile_enable(net, 3): no events; each contraction launch stamps %globaltimer (first CTA in, last CTA out).
    Every launch is seen, the union of the intervals is positive and no longer than their sum, the FLOPs are the
    algorithmic ones, the step computes the same bits, and a second read starts from cleared slots."""
    dims, B, steps = [512, 384, 256, 100], 256, 3
    L = len(dims) - 1
    rng = np.random.RandomState(29)
    x = rng.rand(B, dims[0]).astype(np.float32)
    lab = rng.randint(0, dims[-1], B).astype(np.int32)
    flops_step = 2.0 * B * (3 * sum(a * b for a, b in zip(dims[:-1], dims[1:])) - dims[0] * dims[1])
    ws = []
    for mode in (0, 3):
        net = wide_net(g, dims)
        if mode:
            net.profileEnable(3)
        for _ in range(steps):
            net.gradientStepLabeled(x, lab, 0.05, 0.9)
        if mode:
            busy, tot, n, fl = net.profileReadStamps()
            assert n == steps * (3 * L - 1) and fl == steps * flops_step
            assert 0 < busy <= tot * (1 + 1e-12)
            assert tot / n < 1.0                        # ms per launch of these small contractions: sane clock units
            assert 300.0 < net.profileStampClockMHz() < 3000.0   # CTA 0's cycles / ns: an SM clock
            assert net.profileReadStamps()[2] == 0      # cleared
            net.profileEnable(False)
        ws.append(net.getWeights())
        net.close()
    assert np.array_equal(ws[0], ws[1])
