"""GPU tests of the checkpoint file (b2nn_checkpoint_save / _load, csrc/checkpoint.cu): what gradientStep mutates
(S:327-344: weights, previousUpdate, time) survives a save / load bit for bit, training resumed from the file is
identical to training that never stopped, and a file that does not fit or is damaged leaves the net untouched."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def batch(rng, B, d0, dL):
    x = rng.rand(B, d0)
    lab = rng.randint(0, dL, B).astype(np.int32)
    t = np.zeros((B, dL))
    t[np.arange(B), lab] = 1.0
    return x, t, lab


def test_f64_resume_is_bit_identical(g, tmp_path):
    dims = [20, 16, 12, 5]
    rng = np.random.RandomState(0)
    batches = [batch(rng, 8, 20, 5) for _ in range(5)]
    a = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    for x, t, _ in batches[:3]:
        a.gradientStep((x, t), 0.01, 0.9)
    assert a.getTime() == 3
    path = str(tmp_path / "f64.ckpt")
    a.saveCheckpoint(path)
    info = g.capi.checkpoint_info(path)
    assert info["layerDims"] == dims and info["time"] == 3 and info["numWeights"] == a.numWeights
    assert info["precision"] == g.capi.PREC_F64 and info["kind"] == g.capi.NET_SOFTMAX_CE
    b = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    b.loadCheckpoint(path)
    assert b.getTime() == 3
    assert np.array_equal(a.getWeights(), b.getWeights()) and np.array_equal(a.getVelocity(), b.getVelocity())
    assert a.getVelocity().any()
    for x, t, _ in batches[3:]:
        a.gradientStep((x, t), 0.01, 0.9)
        b.gradientStep((x, t), 0.01, 0.9)
    assert a.getTime() == b.getTime() == 5
    assert np.array_equal(a.getWeights(), b.getWeights()) and np.array_equal(a.getVelocity(), b.getVelocity())


def test_bf16_resume_is_bit_identical(g, tmp_path):
    dims = [256, 128, 32]
    flags = g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX
    mk = lambda: g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                                flags=flags)
    rng = np.random.RandomState(1)
    batches = [batch(rng, 64, 256, 32) for _ in range(4)]
    a = mk()
    for x, _, lab in batches[:2]:
        a.gradientStepLabeled(x.astype(np.float32), lab, 0.05, 0.9)
    path = str(tmp_path / "bf16.ckpt")
    a.saveCheckpoint(path)
    b = mk()
    b.loadCheckpoint(path)
    assert np.array_equal(a.getWeights(), b.getWeights()) and np.array_equal(a.getVelocity(), b.getVelocity())
    for x, _, lab in batches[2:]:
        a.gradientStepLabeled(x.astype(np.float32), lab, 0.05, 0.9)
        b.gradientStepLabeled(x.astype(np.float32), lab, 0.05, 0.9)
    assert np.array_equal(a.getWeights(), b.getWeights()) and np.array_equal(a.getVelocity(), b.getVelocity())
    assert a.getTime() == b.getTime() == 4


def test_misfit_or_damaged_file_leaves_the_net_untouched(g, tmp_path):
    a = g.SoftmaxCrossEntropyNeuralNet([20, 16, 5], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    x, t, _ = batch(np.random.RandomState(2), 8, 20, 5)
    a.gradientStep((x, t), 0.01, 0.9)
    path = str(tmp_path / "a.ckpt")
    a.saveCheckpoint(path)
    other = g.SoftmaxCrossEntropyNeuralNet([20, 12, 5], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    w0 = other.getWeights()
    with pytest.raises(ValueError):
        other.loadCheckpoint(path)                                   # other layerDims
    sig = g.GeneralNeuralNet([20, 16, 5], g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME,
                             g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME)
    with pytest.raises(ValueError):
        sig.loadCheckpoint(path)                                     # other class / functions
    blob = bytearray(open(path, "rb").read())
    blob[len(blob) // 2] ^= 0x40                                     # one flipped bit inside the weight arrays
    bad = str(tmp_path / "bad.ckpt")
    open(bad, "wb").write(bytes(blob))
    same = g.SoftmaxCrossEntropyNeuralNet([20, 16, 5], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    w1 = same.getWeights()
    with pytest.raises(ValueError):
        same.loadCheckpoint(bad)                                     # checksum
    assert np.array_equal(other.getWeights(), w0) and np.array_equal(same.getWeights(), w1)
    assert same.getTime() == 0


def test_f64_checkpoint_resumes_in_bf16(g, tmp_path):
    """the file holds FP64 in the boundary layout, so a parity-mode run can be continued on the tensor cores"""
    dims = [64, 32, 8]
    a = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    x, t, _ = batch(np.random.RandomState(3), 8, 64, 8)
    a.gradientStep((x, t), 0.01, 0.9)
    path = str(tmp_path / "x.ckpt")
    a.saveCheckpoint(path)
    b = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                       flags=g.capi.FLAG_STABLE_SOFTMAX)
    b.loadCheckpoint(path)
    assert np.array_equal(b.getWeights(), a.getWeights().astype(np.float32).astype(np.float64))
    assert np.array_equal(b.getVelocity(), a.getVelocity().astype(np.float32).astype(np.float64))
    assert b.getTime() == 1
