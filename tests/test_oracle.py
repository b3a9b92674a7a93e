"""CPU tests of the oracle itself: pins the java.util.Random restatement to the published
known answers, checks the restated backprop against finite differences of the restated loss, and
checks the harness pieces (sampler, IDX readers, `>=` argmax, log row) against the reference's own
fixtures (label files, logs/trainLog.csv)."""
import json
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def kat(golden_dir):
    with open(os.path.join(golden_dir, "java_random_kat.json")) as f:
        return json.load(f)


def test_jrandom_known_answers(orc, kat):
    assert orc.JavaRandom(1).next_int() == kat["next_int"]
    r = orc.JavaRandom(1)
    assert [r.next_double() for _ in range(4)] == kat["next_double_x4"]
    r = orc.JavaRandom(1)
    assert [r.next_gaussian(), r.next_gaussian()] == kat["next_gaussian_x2"]
    r = orc.JavaRandom(1)
    assert [r.next_int(b) for b in kat["sampler_bounds"]] == kat["sampler_draws"]


def test_jrandom_power_of_two_bound(orc):
    # nextInt(bound) power-of-two shortcut: (bound * next(31)) >> 31
    r1, r2 = orc.JavaRandom(7), orc.JavaRandom(7)
    for _ in range(100):
        v = r1.next_int(16)
        raw = r2.next_int()  # next(32); next(31) is its top 31 bits
        assert v == (((raw & 0xFFFFFFFF) >> 1) * 16) >> 31


@pytest.mark.parametrize("key,dims", [("net_784_300_100_10", [784, 300, 100, 10]),
                                       ("net_784_100_50_10", [784, 100, 50, 10])])
def test_weight_init_known_answers(orc, kat, key, dims):
    n = orc.OracleNet(dims)
    w = n.get_weights()
    assert w.size == kat[key]["n_weights"]
    s = 0.0
    for v in w:
        s += v
    assert s == kat[key]["sequential_sum"]
    assert w[-1] == kat[key]["last_weight"]
    assert hex(n.rng_state) == kat[key]["lcg_state_after"]
    if key == "net_784_300_100_10":
        assert list(w[:4]) == kat["first_weights_x4"]
    assert w.min() >= -0.5 and w.max() < 0.5
    assert not n.get_velocity().any()


def test_general_net_same_init_stream(orc):
    a = orc.OracleNet([30, 20, 10])
    b = orc.OracleNet([30, 20, 10], kind=orc.NET_GENERAL, inner_act="sigmoid", last_act="sigmoid")
    assert (a.get_weights() == b.get_weights()).all()  # both use Random(1) (G:130)


def _fd_grad(net, x, t, eps=1e-6):
    w0 = net.get_weights()
    num = np.zeros_like(w0)
    for q in range(w0.size):
        w = w0.copy()
        w[q] += eps
        net.set_weights(w)
        lp = net.loss(x, t)
        w[q] -= 2 * eps
        net.set_weights(w)
        lm = net.loss(x, t)
        num[q] = (lp - lm) / (2 * eps)
    net.set_weights(w0)
    return num


@pytest.mark.parametrize("cfg", [
    dict(kind=0, inner_act="leaky_relu", inner_param=0.01),
    dict(kind=0, inner_act="sigmoid", inner_param=0.0),
    dict(kind=1, inner_act="sigmoid", inner_param=0.0, last_act="sigmoid"),
    dict(kind=1, inner_act="tanh", inner_param=0.0, last_act="identity"),
])
def test_gradient_matches_finite_differences(orc, cfg):
    rng = np.random.RandomState(3)
    net = orc.OracleNet([9, 7, 5, 4], **cfg)
    x = rng.rand(9) - 0.3
    t = np.zeros(4)
    t[2] = 1.0
    g = net.weight_gradient(x, t)
    num = _fd_grad(net, x, t)
    assert np.abs(g - num).max() < 1e-7


def test_input_layer_is_activated(orc):
    # Q2: a0 = f(x); with sigmoid inner activation the first-layer sums use sigmoid(x)
    net = orc.OracleNet([3, 2], kind=orc.NET_GENERAL, inner_act="sigmoid", last_act="identity")
    W = net.get_weights().reshape(3, 2)
    x = np.array([0.0, 1.0, -2.0])
    y = net.propagate(x)
    a0 = 1.0 / (1.0 + np.exp(-x))
    assert np.allclose(y, a0 @ W, rtol=0, atol=1e-15)


def test_softmax_not_max_normalised_overflows(orc):
    # Q4: the reference softmax never subtracts the max -> NaN once a logit exceeds ~709
    net = orc.OracleNet([1, 2], inner_act="identity")
    net.set_weights(np.array([800.0, 1.0]))
    with np.errstate(all="ignore"):
        y = net.propagate(np.array([1.0]))
    assert np.isnan(y[0])


def test_gradient_step_faithful_equals_streamlined(orc):
    rng = np.random.RandomState(0)
    x = rng.rand(6, 12)
    t = np.zeros((6, 4))
    t[np.arange(6), rng.randint(0, 4, 6)] = 1
    a, b = orc.OracleNet([12, 8, 4]), orc.OracleNet([12, 8, 4])
    for _ in range(5):
        a.gradient_step(x, t, 0.05, 0.9, faithful=True)
        b.gradient_step(x, t, 0.05, 0.9, faithful=False)
    assert (a.get_weights() == b.get_weights()).all()
    assert (a.get_velocity() == b.get_velocity()).all()


def test_gradient_step_is_momentum_sgd(orc):
    # Q7: v <- step*G/B + mu*v ; W <- W - v
    rng = np.random.RandomState(1)
    x = rng.rand(4, 6)
    t = np.zeros((4, 3))
    t[np.arange(4), [0, 1, 2, 1]] = 1
    net = orc.OracleNet([6, 5, 3])
    w0 = net.get_weights()
    G = sum(net.weight_gradient(x[i], t[i]) for i in range(4))
    net.gradient_step(x, t, 0.1, 0.9)
    v1 = net.get_velocity()
    assert np.allclose(v1, 0.1 * G / 4, rtol=1e-15, atol=0)
    assert np.allclose(net.get_weights(), w0 - v1, rtol=1e-15, atol=0)
    w1 = net.get_weights()
    G2 = sum(net.weight_gradient(x[i], t[i]) for i in range(4))
    net.gradient_step(x, t, 0.1, 0.9)
    assert np.allclose(net.get_velocity(), 0.1 * G2 / 4 + 0.9 * v1, rtol=1e-14, atol=0)
    assert np.allclose(net.get_weights(), w1 - net.get_velocity(), rtol=1e-15, atol=0)


def test_noise_branch_yields_nan_like_the_reference(orc):
    # Q8: sqrt(gaussian * step / ...) is NaN for negative draws
    net = orc.OracleNet([4, 3])
    x = np.ones((1, 4))
    t = np.array([[1.0, 0, 0]])
    with np.errstate(all="ignore"):
        net.gradient_step(x, t, 0.1, 0.9, noise=True)
    w = net.get_weights()
    assert np.isnan(w).any() and not np.isnan(w).all()


def test_sampler_without_replacement_and_refill(orc):
    s = orc.Sampler(50)
    seen = np.concatenate([s.sample(5) for _ in range(10)])
    assert sorted(seen) == list(range(50))           # one epoch covers every row once
    again = np.concatenate([s.sample(5) for _ in range(10)])
    assert sorted(again) == list(range(50))          # refill starts a new epoch


def test_argmax_last_max_wins(orc):
    # Q10: `>=` => the LAST maximal index wins ties
    net = orc.OracleNet([2, 3], inner_act="identity")
    net.set_weights(np.zeros(6))                     # all outputs equal -> argmax must be 2
    hits, pred = net.evaluate(np.ones((4, 2)), np.array([2, 2, 0, 1]))
    assert list(pred) == [2, 2, 2, 2] and hits == 2


def test_idx_label_files(orc, golden_dir):
    tr = orc.read_idx_labels(os.path.join(golden_dir, "mnist", "train-labels-idx1-ubyte"))
    te = orc.read_idx_labels(os.path.join(golden_dir, "mnist", "t10k-labels-idx1-ubyte"))
    assert tr.size == 60000 and te.size == 10000
    assert list(np.bincount(tr)) == [5923, 6742, 5958, 6131, 5842, 5421, 5918, 6265, 5851, 5949]
    assert list(np.bincount(te)) == [980, 1135, 1032, 1010, 982, 892, 958, 1028, 974, 1009]


def test_idx_images_roundtrip(orc, tmp_path):
    import struct
    raw = np.arange(2 * 28 * 28, dtype=np.uint32).astype(np.uint8).reshape(2, 784)
    p = tmp_path / "img"
    with open(p, "wb") as f:
        f.write(struct.pack(">iiii", 2051, 2, 28, 28))
        f.write(raw.tobytes())
    x = orc.read_idx_images(str(p))
    assert x.shape == (2, 784)
    assert (x == raw / 255.0).all()
    with open(p, "r+b") as f:
        f.write(struct.pack(">i", 2049))             # wrong magic must be rejected (Q21)
    with pytest.raises(IOError):
        orc.read_idx_images(str(p))


def test_log_row_matches_reference_log(orc, golden_dir):
    # the five rows the reference recorded (logs/trainLog.csv) re-format to themselves
    with open(os.path.join(golden_dir, "trainLog_reference.csv")) as f:
        rows = [r for r in f.read().splitlines() if r]
    assert len(rows) == 5
    for r in rows:
        dims, it, step, b, mom, noise, tra, tea = r.split(",")
        out = orc.format_log_row([int(d) for d in dims.split("-")], int(it), float(step), int(b), float(mom),
                                 noise == "true", float(tra), float(tea))
        assert out == r + "\n"


def test_train_matches_manual_loop(orc):
    rng = np.random.RandomState(5)
    x = rng.rand(40, 10)
    lab = rng.randint(0, 3, 40)
    t = np.zeros((40, 3))
    t[np.arange(40), lab] = 1
    a, b = orc.OracleNet([10, 6, 3]), orc.OracleNet([10, 6, 3])
    a.train(x, t, 12, 4, 0.05, 0.9)                 # internal sampler(1)
    s = orc.Sampler(40)
    for _ in range(12):
        i = s.sample(4)
        b.gradient_step(x[i], t[i], 0.05, 0.9)
    assert (a.get_weights() == b.get_weights()).all()
    assert abs(a.validate(x, t, 7) - np.mean([a.loss(x[i], t[i]) for i in range(7)])) < 1e-15


def test_worker_threads_are_bit_identical(orc):
    """orc.set_threads(T): per-sample gradients on T threads, summed in batch order -- the same bits as the serial
    loops (gradientStep S:305-342, validate T:102-113, the accuracy pass M:159-172) for both net classes"""
    rng = np.random.RandomState(5)
    for kind_kw, dims, B in (({}, [30, 17, 11, 6], 8),
                             (dict(kind=orc.NET_GENERAL, inner_act="sigmoid", inner_param=0.0, last_act="sigmoid"),
                              [30, 12, 9, 5], 16)):
        n_rows = 96
        x = rng.rand(n_rows, dims[0])
        lab = rng.randint(0, dims[-1], n_rows).astype(np.int32)
        t = np.zeros((n_rows, dims[-1]))
        t[np.arange(n_rows), lab] = 1.0
        out = []
        for threads in (1, 4):
            orc.set_threads(threads)
            net = orc.OracleNet(dims, **kind_kw)
            losses = net.train(x, t, 12, B, 0.05, 0.9, want_loss=True)
            out.append((net.get_weights(), net.get_velocity(), losses, net.validate(x, t, 50), net.evaluate(x, lab)))
        orc.set_threads(1)
        a, b = out
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        assert a[3] == b[3] and a[4][0] == b[4][0] and np.array_equal(a[4][1], b[4][1])
