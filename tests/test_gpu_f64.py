"""GPU parity tests of the FP64 path against the CPU oracle, through the C-ABI.

Tolerances (stated here): the kernels keep the reference's evaluation order and do not
contract multiply-add, so only exp()/log() (each <= 1 ulp different from libm) separate the two:
  - initial weights: bit-exact;
  - propagate / loss / one-sample gradient: |diff| <= 1e-13 * max(1, |ref|);
  - weights after 1 step: <= 1e-13 * max(1,|W|); after 300 steps: relative <= 1e-9;
  - per-step mean loss: relative <= 1e-9;
  - argmax over the evaluation set: identical.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CONFIGS = {
    "C1": dict(dims=[784, 100, 50, 10], B=8, step=0.01, kind="softmax"),      # logs/trainLog.csv:3
    "C2": dict(dims=[784, 300, 100, 10], B=16, step=0.0125, kind="softmax"),  # MNISTTrainer.java:224-230
    "C3": dict(dims=[784, 100, 50, 10], B=8, step=0.01, kind="general"),      # sigmoid + half squared error
}


def make_nets(g, orc, cfg, dims=None):
    dims = dims or cfg["dims"]
    if cfg["kind"] == "softmax":
        net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
        ref = orc.OracleNet(dims)
    else:
        net = g.GeneralNeuralNet(dims, g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME,
                                 g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME)
        ref = orc.OracleNet(dims, kind=orc.NET_GENERAL, inner_act="sigmoid", inner_param=0.0, last_act="sigmoid")
    return net, ref


def make_data(n, d0, dL, seed):
    rng = np.random.RandomState(seed)
    x = np.round(rng.rand(n, d0) * 255) / 255.0   # pixel-like
    x[rng.rand(n, d0) < 0.5] = 0.0
    lab = rng.randint(0, dL, n).astype(np.int32)
    t = np.zeros((n, dL))
    t[np.arange(n), lab] = 1.0
    return x, t, lab


def close(a, b, tol):
    a, b = np.asarray(a), np.asarray(b)
    return np.all(np.abs(a - b) <= tol * np.maximum(1.0, np.abs(b)))


@pytest.mark.parametrize("name", ["C1", "C2", "C3"])
def test_init_bit_exact_and_dims(g, orc, name):
    net, ref = make_nets(g, orc, CONFIGS[name])
    assert (net.getWeights() == ref.get_weights()).all()
    assert not net.getVelocity().any()
    assert net.getInputDim() == 784 and net.getOutputDim() == 10
    assert net.numWeights == ref.n_weights


@pytest.mark.parametrize("name", ["C1", "C2", "C3"])
def test_propagate_loss_gradient(g, orc, name):
    cfg = CONFIGS[name]
    net, ref = make_nets(g, orc, cfg)
    # the reference init gives huge logits on 784 inputs; damp the weights so softmax stays finite
    w = ref.get_weights() * 0.05
    net.setWeights(w)
    ref.set_weights(w)
    x, t, _ = make_data(37, 784, 10, 1)
    y = net.propagate(x)
    yr = ref.propagate(x)
    assert close(y, yr, 1e-13)
    assert close(net.propagate(x[0]), yr[0], 1e-13)
    l = net.calculateLoss(x, t)
    lr = np.array([ref.loss(x[i], t[i]) for i in range(37)])
    assert close(l, lr, 1e-13)
    gr = ref.weight_gradient(x[3], t[3])
    gg = net.calculateWeightGradient(x[3], t[3])
    flat = np.concatenate([gg[l].ravel() for l in range(3)])
    assert gg[0].shape == (784, cfg["dims"][1]) and gg[2].shape == (cfg["dims"][2], 10)
    assert close(flat, gr, 1e-13)


@pytest.mark.parametrize("name", ["C1", "C2", "C3"])
def test_gradient_steps(g, orc, name):
    cfg = CONFIGS[name]
    net, ref = make_nets(g, orc, cfg)
    w = ref.get_weights() * 0.05
    net.setWeights(w)
    ref.set_weights(w)
    B = cfg["B"]
    x, t, _ = make_data(B * 10, 784, 10, 2)
    loss = net.gradientStep((x[:B], t[:B]), cfg["step"], 0.9, returnLoss=True)
    ref_loss = np.mean([ref.loss(x[i], t[i]) for i in range(B)])
    ref.gradient_step(x[:B], t[:B], cfg["step"], 0.9)
    assert abs(loss - ref_loss) <= 1e-12 * max(1.0, abs(ref_loss))
    assert close(net.getWeights(), ref.get_weights(), 1e-13)
    assert close(net.getVelocity(), ref.get_velocity(), 1e-13)
    for k in range(1, 10):
        pairs = [(x[k * B + i], t[k * B + i]) for i in range(B)]   # the Map<double[],double[]> form
        net.gradientStep(pairs, cfg["step"], 0.9)
        ref.gradient_step(x[k * B:(k + 1) * B], t[k * B:(k + 1) * B], cfg["step"], 0.9)
    assert close(net.getWeights(), ref.get_weights(), 1e-12)


@pytest.mark.parametrize("name", ["C1", "C2", "C3"])
def test_trainer_indexed_run_and_eval(g, orc, name):
    """NeuralNetTrainer.train over a dataset: 300 steps with the reference sampler's indices."""
    cfg = CONFIGS[name]
    net, ref = make_nets(g, orc, cfg)
    w = ref.get_weights() * 0.05
    net.setWeights(w)
    ref.set_weights(w)
    n, B, iters = 2000, cfg["B"], 300
    x, t, lab = make_data(n, 784, 10, 3)
    tr = g.NeuralNetTrainer((x, t), net)
    idx = g.Sampler(n, seed=1).draw(iters, B)
    l0 = net.kernelLaunches()
    losses = tr.trainIndexed(idx, cfg["step"], 0.9, wantLoss=True)
    assert net.kernelLaunches() - l0 == 1          # the whole run is ONE persistent kernel (K1), no fallback
    ref_losses = ref.train(x, t, iters, B, cfg["step"], 0.9, idx=idx, want_loss=True)
    assert np.all(np.abs(losses - ref_losses) <= 1e-9 * np.maximum(1.0, np.abs(ref_losses)))
    wg, wr = net.getWeights(), ref.get_weights()
    assert np.max(np.abs(wg - wr) / np.maximum(1e-3, np.abs(wr))) <= 1e-9
    # validate(601)-style mean loss and `>=` argmax accuracy
    v = tr.validate(n // 100 + 1)
    assert abs(v - ref.validate(x, t, n // 100 + 1)) <= 1e-9 * max(1.0, abs(v))
    hits, mean_loss, pred = tr.dataMaster.evaluate(want_pred=True)
    rhits, rpred = ref.evaluate(x, lab)
    assert (pred == rpred).all() and hits == rhits


def test_trainer_train_equals_oracle_train(g, orc):
    """train() with its own Random(1) sampler == oracle's train() with its own sampler."""
    cfg = CONFIGS["C1"]
    dims = [60, 30, 20, 10]
    net, ref = make_nets(g, orc, cfg, dims)
    x, t, _ = make_data(500, 60, 10, 4)
    tr = g.NeuralNetTrainer([(x[i], t[i]) for i in range(500)], net)
    tr.train(150, 0.01, 8, 0.9, False, None, None)
    ref.train(x, t, 150, 8, 0.01, 0.9)
    assert close(net.getWeights(), ref.get_weights(), 1e-10)


def test_observer_stream_format(g, orc, tmp_path):
    import io
    dims = [20, 10, 5]
    net, ref = make_nets(g, orc, CONFIGS["C1"], dims)
    x, t, _ = make_data(300, 20, 5, 5)
    tr = g.NeuralNetTrainer((x, t), net)
    buf = io.StringIO()
    tr.train(5, 0.01, 4, 0.9, False, None, buf)
    lines = buf.getvalue().splitlines()
    s = orc.Sampler(300)
    for i in range(5):
        b = s.sample(4)
        ref.gradient_step(x[b], t[b], 0.01, 0.9)
        assert lines[i] == "%d,%.2f" % (i, ref.validate(x, t, 300 // 100 + 1))   # T:71


def test_edge_cases_and_errors(g, orc):
    dims = [5, 3, 2]
    net, ref = make_nets(g, orc, CONFIGS["C1"], dims)
    x, t, _ = make_data(3, 5, 2, 6)
    net.gradientStep((x[:1], t[:1]), 0.1, 0.0)            # B = 1, no momentum
    ref.gradient_step(x[:1], t[:1], 0.1, 0.0)
    assert close(net.getWeights(), ref.get_weights(), 1e-13)
    with pytest.raises(NotImplementedError):
        net.gradientStep((x, t), 0.1, 0.9, noise=True)    # Q8
    with pytest.raises(ValueError):
        net.gradientStep([], 0.1, 0.9)                    # S:300
    with pytest.raises(ValueError):
        net.gradientStep((x, t), 0.0, 0.9)                # S:301
    with pytest.raises(ValueError):
        net.propagate(np.zeros(4))                        # S:166
    with pytest.raises(ValueError):
        g.NeuralNetTrainer((x, t), net).train(1, 0.1, 3, 0.9)   # batchSize < size strict (T:63)
    # all-equal outputs: `>=` argmax picks the last index (Q10)
    net.setWeights(np.zeros(net.numWeights))
    ds = g.Dataset(net, x, labels=np.array([1, 1, 0], dtype=np.int32))
    hits, _, pred = ds.evaluate(want_pred=True)
    assert list(pred) == [1, 1, 1] and hits == 2


def test_two_layer_and_deep_ragged(g, orc):
    for dims in ([7, 3], [33, 17, 65, 9, 4]):
        net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
        ref = orc.OracleNet(dims)
        w = ref.get_weights() * 0.3
        net.setWeights(w)
        ref.set_weights(w)
        x, t, _ = make_data(70, dims[0], dims[-1], 7)     # M > 32 exercises the 64-row tile path
        for _ in range(3):
            net.gradientStep((x, t), 0.05, 0.9)
            ref.gradient_step(x, t, 0.05, 0.9)
        assert close(net.getWeights(), ref.get_weights(), 1e-12)
        assert close(net.propagate(x), ref.propagate(x), 1e-12)


@pytest.mark.parametrize("kind", ["softmax", "general"])
def test_fused_persistent_kernel_vs_general_path_and_oracle(g, orc, kind, monkeypatch):
    """K1 (one persistent cooperative kernel for the whole run) against the general multi-kernel
    path (B2NN_F64_FUSED=0) and the oracle, on non-pixel inputs and ragged dims / batch."""
    dims, B, iters, n = [33, 47, 21, 7], 5, 60, 400
    cfg = dict(kind=kind)
    rng = np.random.RandomState(11)
    x = rng.randn(n, 33) * 0.7                      # not k/255: exercises the FP64-row loader
    lab = rng.randint(0, 7, n).astype(np.int32)
    t = np.zeros((n, 7))
    t[np.arange(n), lab] = 1.0
    idx = g.Sampler(n, seed=1).draw(iters, B)
    results = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("B2NN_F64_FUSED", mode)
        net, ref = make_nets(g, orc, cfg, dims)
        tr = g.NeuralNetTrainer((x, t), net)
        losses = tr.trainIndexed(idx, 0.02, 0.9, wantLoss=True)
        results[mode] = (net.getWeights(), net.getVelocity(), losses, net.kernelLaunches())
    ref_losses = ref.train(x, t, iters, B, 0.02, 0.9, idx=idx, want_loss=True)
    wr, vr = ref.get_weights(), ref.get_velocity()
    for mode in ("1", "0"):
        w, v, losses, _ = results[mode]
        assert np.all(np.abs(losses - ref_losses) <= 1e-10 * np.maximum(1.0, np.abs(ref_losses))), mode
        assert close(w, wr, 1e-10) and close(v, vr, 1e-10), mode
    assert results["1"][3] < results["0"][3] / 20    # the fused run is a handful of launches


def test_fused_kernel_pixel_bytes_equal_double_rows(g, orc, monkeypatch):
    """rows that are exactly byte/255.0 are staged as bytes for K1; the result must be identical to
    feeding the same doubles through the FP64-row loader (a perturbed copy defeats the detection)."""
    dims, B, iters, n = [784, 100, 50, 10], 8, 40, 600
    x, t, lab = make_data(n, 784, 10, 13)
    idx = g.Sampler(n, seed=1).draw(iters, B)
    out = []
    for perturb in (False, True):
        net, _ = make_nets(g, orc, CONFIGS["C1"], dims)
        xx = x.copy()
        if perturb:
            xx[-1, -1] = 0.123456789                # row n-1 is never drawn below
        idx2 = np.where(idx == n - 1, 0, idx).astype(np.int32)
        tr = g.NeuralNetTrainer((xx, t), net)
        tr.trainIndexed(idx2, 0.01, 0.9)
        out.append(net.getWeights())
    assert np.array_equal(out[0], out[1])
