"""Full-size checks of the tensor-core path (BASELINE configs[3]: 4096-4096-4096-1000, batch 4096; configs[4]:
8192-wide layers) through the C-ABI.  The FP64 oracle needs minutes per sample at this width (bench.py
cpu_baseline), so at full size the comparison partner is

  * for a BARE contraction: the exact product of the operands rounded the way the tensor core reads them
    (BF16 / TF32), accumulated in FP64 by numpy -- the three operand major-ness combinations of a training step
    (forward K-major x MN-major, delta K-major x K-major, gradient MN-major x MN-major with K = batch), each isolated
    by choosing weights / inputs that make the other contractions of the step exact (zeros, identity matrix);
  * for a whole gradientStep (S:297-346): the numpy rounding models of tests/test_gpu_wide.py / test_gpu_z_tf32.py
    (validated against the oracle at widths the oracle can run) evaluated at full size;
  * size-independent properties: with momentum 0 the update of the full batch is the mean of the updates of its two
    halves (the batch sum S:305-322 divided by batch.size() S:333), and a step is deterministic.
4096^3 on 74 CTA pairs is 3.46 waves, so these shapes are also the only ones that reach the half-tile tail wave with
every pair busy."""
import numpy as np
import pytest

from tests.test_gpu_wide import bf16_round, emulate_bf16_step
from tests.test_gpu_z_tf32 import emulate_tf32_step, tf32_round

pytestmark = pytest.mark.gpu

DIMS, B = [4096, 4096, 4096, 1000], 4096
ROUND = {"bf16": bf16_round, "tf32": tf32_round}
# bare contraction against exact-operand FP64: what is left is the FP32 accumulation order
BARE_TOL = {"bf16": 1e-3, "tf32": 5e-5}


def prec_id(g, prec):
    return g.capi.PREC_BF16 if prec == "bf16" else g.capi.PREC_TF32


def make(g, precision, dims=DIMS):
    return g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=precision,
                                          flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)


def identity_net(g, dims, precision):
    """GeneralNeuralNet with identity activations everywhere and half squared error: every epilogue is a plain store"""
    return g.GeneralNeuralNet(dims, g.IDENTITY, g.IDENTITY_PRIME, g.IDENTITY, g.IDENTITY_PRIME,
                              g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME, precision=precision)


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
@pytest.mark.parametrize("M,K,N", [(4096, 4096, 4096), (2048, 8192, 8192)])
def test_bare_forward_contraction_full_size(g, prec, M, K, N):
    """y = x W (S:187-192): A K-major, B MN-major, FP32 store epilogue"""
    rnd = ROUND[prec]
    net = identity_net(g, [K, N], prec_id(g, prec))
    rng = np.random.RandomState(K + N)
    W = rnd(rng.randn(K, N) * 0.05)
    x = rnd(rng.randn(M, K))
    net.setWeights(W.ravel())
    y = net.propagate(x)
    ref = x @ W
    err = np.abs(y - ref).max()
    assert err <= BARE_TOL[prec] * np.abs(ref).max(), (prec, err, np.abs(ref).max())


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
@pytest.mark.parametrize("Bt,K,N", [(4096, 4096, 4096), (1024, 8192, 8192)])
def test_bare_gradient_contraction_full_size(g, prec, Bt, K, N):
    """G = A^T D (S:279-283 + S:305-317): both operands MN-major, K = batch.  With W = 0 the output is 0, so the
    output delta is exactly -t (S: loss' = a - t, identity f' = 1), and with step 1 and a power-of-two batch the stored
    velocity is G / B without rounding."""
    rnd = ROUND[prec]
    net = identity_net(g, [K, N], prec_id(g, prec))
    net.setWeights(np.zeros(net.numWeights))
    rng = np.random.RandomState(Bt + N)
    x = rnd(rng.randn(Bt, K))
    t = rnd(rng.randn(Bt, N))
    net.gradientStep((x, t), 1.0, 0.0)
    v = net.getVelocity().reshape(K, N) * Bt
    ref = -(x.T @ t)
    err = np.abs(v - ref).max()
    assert err <= BARE_TOL[prec] * np.abs(ref).max(), (prec, err, np.abs(ref).max())
    assert np.array_equal(net.getWeights().reshape(K, N) * Bt, -v)          # W -= v from zero (S:338)


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_bare_delta_contraction_full_size(g, prec):
    """D_1 = D_2 W_1^T (S:272-278): both operands K-major.  W_0 = 0 makes a_1 = 0, the output 0 and D_2 = -t; the batch
    is the 4096 x 4096 identity matrix, so the layer-0 gradient A_0^T D_1 is D_1 itself (one non-zero product per sum)
    and the stored velocity of W_0 is D_1 / B."""
    n = 4096
    rnd = ROUND[prec]
    net = identity_net(g, [n, n, n], prec_id(g, prec))
    rng = np.random.RandomState(77)
    W1 = rnd(rng.randn(n, n) * 0.05)
    t = rnd(rng.randn(n, n))
    net.setWeights(np.concatenate([np.zeros(n * n), W1.ravel()]))
    net.gradientStep((np.eye(n), t), 1.0, 0.0)
    v = net.getVelocity()
    d1 = v[:n * n].reshape(n, n) * n
    ref = -(t @ W1.T)
    # D_1 reaches the velocity through the layer-0 gradient contraction, which reads it as an operand: BF16 nets store it
    # rounded to BF16 (relative 2^-9), TF32 nets store FP32 but TMA rounds it to TF32 (relative 2^-11) on the way in
    slack = (2.0 ** -8 if prec == "bf16" else 2.0 ** -10) * np.abs(ref)
    assert np.all(np.abs(d1 - ref) <= BARE_TOL[prec] * np.abs(ref).max() + slack), np.abs(d1 - ref).max()
    assert not v[n * n:].any()                                              # a_1 = 0: no gradient for W_1


def report(line):
    """measured errors, kept when B2NN_TEST_REPORT names a file (profiles/r02_fullsize_parity_report.txt)"""
    import os
    if os.environ.get("B2NN_TEST_REPORT"):
        with open(os.environ["B2NN_TEST_REPORT"], "a") as f:
            f.write(line + "\n")


# Per-layer bound of one step against the rounding model.  With a SMOOTH activation (sigmoid) the model and the kernels
# differ only by FP32 accumulation order and 1-ulp effects: 1e-2 (BF16) / 1e-3 (TF32) in every layer, the bounds of the
# small-size tests (measured at C4 size: <= 4.9e-4 / <= 9e-5).  With leaky-ReLU a unit whose pre-activation lies within
# the FP32 accumulation noise of zero takes the other slope in one of the two computations: its delta is off by ~100 %,
# and every gradient BELOW it inherits the difference (Frobenius ~ sqrt(fraction of flipped units); measured at C4 size:
# 4.1e-3 BF16, 3.4e-3 TF32 -- the same in both precisions, as a flip count should be); the top layer has no f' above it
# and keeps the tight bound (measured 3.4e-4 / 1.2e-4).  Measured values: profiles/r02_fullsize_parity_report.txt.
STEP_TOL = {("bf16", "sigmoid"): (1e-2, 1e-2), ("tf32", "sigmoid"): (1e-3, 1e-3),
            ("bf16", "leaky"): (1e-2, 1e-2), ("tf32", "leaky"): (1e-3, 1e-2)}   # (top layer, layers below)


def check_step_vs_model(tag, prec, act, dims, v, emu_v, lower_tol_scale=1.0):
    top, lower = STEP_TOL[(prec, act)]
    L, off = len(dims) - 1, 0
    for l in range(L):
        n = dims[l] * dims[l + 1]
        a, e = v[off:off + n], emu_v[l].ravel()
        rel = np.linalg.norm(a - e) / np.linalg.norm(e)
        report("%s %s %s layer %d rel %.3g" % (tag, prec, act, l, rel))
        assert rel <= (top if l == L - 1 else lower * lower_tol_scale), (tag, prec, act, l, rel)
        off += n


@pytest.mark.parametrize("act", ["sigmoid", "leaky"])
@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_c4_full_step_vs_rounding_model(g, prec, act):
    """one gradientStep of the real C4 (and of its sigmoid twin, which has no slope discontinuity) against the numpy
    model that rounds where the kernels round, evaluated at full size; batch loss within 1e-3 (BF16) / 1e-4 (TF32)"""
    ltol, emu = {"bf16": (1e-3, emulate_bf16_step), "tf32": (1e-4, emulate_tf32_step)}[prec]
    rng = np.random.RandomState(8)
    x = rng.rand(B, DIMS[0]).astype(np.float32)
    lab = rng.randint(0, DIMS[-1], B).astype(np.int32)
    f, fp = (g.SIGMOID, g.SIGMOID_PRIME) if act == "sigmoid" else (g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    net = g.SoftmaxCrossEntropyNeuralNet(DIMS, f, fp, precision=prec_id(g, prec),
                                         flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
    w0 = net.getWeights()
    loss = net.gradientStepLabeled(x, lab, 0.01, 0.0, returnLoss=True)
    v = net.getVelocity()
    net.close()
    emu_loss, emu_v = emu(DIMS, w0, x, lab, 0.01, act=act)
    report("c4 %s %s loss rel %.3g" % (prec, act, abs(loss - emu_loss) / max(1.0, abs(emu_loss))))
    assert abs(loss - emu_loss) <= ltol * max(1.0, abs(emu_loss)), (loss, emu_loss)
    check_step_vs_model("c4", prec, act, DIMS, v, emu_v)


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_six_matrix_net_vs_oracle(g, orc, prec):
    """BASELINE configs[4]'s shape (7 dims = 6 weight matrices, 1000 classes) at a width the oracle can run: first step
    against the rounding model and the FP64 oracle (loose: leaky-ReLU slope flips compound through five hidden layers,
    see STEP_TOL; twice the C4 allowance below the top layer, measured <= 1e-2), then two more momentum steps with the
    batch loss tracked against the oracle's.  Measured against the oracle: BF16 <= 7.4e-2 (cosine >= 0.9973), TF32
    <= 3.1e-2 (cosine >= 0.9995) below the top layer, 5e-3 / 6e-4 in the top layer."""
    dims, Bs = [384] * 6 + [1000], 128
    otol, cos_min, emu = {"bf16": (2e-1, 0.98, emulate_bf16_step), "tf32": (5e-2, 0.998, emulate_tf32_step)}[prec]
    net = make(g, prec_id(g, prec), dims)
    ref = orc.OracleNet(dims)
    ref.set_weights(net.getWeights())
    rng = np.random.RandomState(21)
    x = rng.rand(Bs, dims[0])
    lab = rng.randint(0, dims[-1], Bs).astype(np.int32)
    t = np.zeros((Bs, dims[-1]))
    t[np.arange(Bs), lab] = 1.0
    w0 = ref.get_weights()
    emu_loss, emu_v = emu(dims, w0, x, lab, 0.05)
    for it in range(3):
        ref_loss = np.mean([ref.loss(x[i], t[i]) for i in range(Bs)])
        loss = net.gradientStepLabeled(x.astype(np.float32), lab, 0.05, 0.9, returnLoss=True)
        ref.gradient_step(x, t, 0.05, 0.9)
        assert abs(loss - ref_loss) <= 2e-2 * max(1.0, abs(ref_loss)), (it, loss, ref_loss)
        if it == 0:
            assert abs(loss - emu_loss) <= 1e-3 * max(1.0, abs(emu_loss))
            du, du_ref = w0 - net.getWeights(), w0 - ref.get_weights()
            check_step_vs_model("six", prec, "leaky", dims, du, emu_v, lower_tol_scale=2.0)
            off = 0
            for l in range(len(dims) - 1):
                n = dims[l] * dims[l + 1]
                a, r = du[off:off + n], du_ref[off:off + n]
                rel = np.linalg.norm(a - r) / np.linalg.norm(r)
                report("six %s vs oracle layer %d rel %.3g cos %.6f" % (prec, l, rel, a @ r / np.linalg.norm(a) / np.linalg.norm(r)))
                assert rel <= otol, (l, rel)
                assert a @ r >= cos_min * np.linalg.norm(a) * np.linalg.norm(r)
                off += n


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_c4_full_batch_update_is_the_mean_of_its_halves(g, prec):
    precision = prec_id(g, prec)
    rng = np.random.RandomState(4)
    x = rng.rand(B, DIMS[0]).astype(np.float32)
    lab = rng.randint(0, DIMS[-1], B).astype(np.int32)
    full = make(g, precision)
    loss_full = full.gradientStepLabeled(x, lab, 0.01, 0.0, returnLoss=True)
    v_full = full.getVelocity()
    full.close()
    v_half, loss_half = [], []
    for lo in (0, B // 2):
        net = make(g, precision)
        loss_half.append(net.gradientStepLabeled(x[lo:lo + B // 2], lab[lo:lo + B // 2], 0.01, 0.0, returnLoss=True))
        v_half.append(net.getVelocity())
        net.close()
    mean = 0.5 * (v_half[0] + v_half[1])
    assert abs(loss_full - 0.5 * (loss_half[0] + loss_half[1])) <= 1e-6 * max(1.0, abs(loss_full))
    off = 0
    for l in range(len(DIMS) - 1):
        n = DIMS[l] * DIMS[l + 1]
        a, m = v_full[off:off + n], mean[off:off + n]
        assert np.linalg.norm(m) > 0
        rel = np.linalg.norm(a - m) / np.linalg.norm(m)
        assert rel <= 1e-3, (prec, l, rel)
        off += n


def test_c4_step_is_deterministic(g):
    rng = np.random.RandomState(5)
    x = rng.rand(B, DIMS[0]).astype(np.float32)
    lab = rng.randint(0, DIMS[-1], B).astype(np.int32)
    out = []
    for _ in range(2):
        net = make(g, g.capi.PREC_BF16)
        for _ in range(2):
            net.gradientStepLabeled(x, lab, 0.01, 0.9)
        out.append(net.getWeights())
        net.close()
    assert np.array_equal(out[0], out[1])
