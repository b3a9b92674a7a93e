"""Full-size (BASELINE configs[3]: 4096-4096-4096-1000, batch 4096) checks of the tensor-core path through
size-independent properties -- the FP64 oracle needs minutes per sample at this width (bench.py cpu_baseline), so
parity at full size rests on what the reference's definition of the batch gradient implies:
  * gradientStep sums per-sample gradients over the batch (S:305-322) and divides by batch.size() (S:333), so with
    momentum 0 the update of the full batch is the mean of the updates of its two halves;  per-row quantities do
    not depend on which other rows share the batch, so only the FP32 summation order of the gradient contraction
    (K = 4096 in one accumulator against two K = 2048 accumulators averaged in FP64) differs:  <= 1e-3 relative
    (Frobenius, per layer; the same bound family as the 2e-3 of the cross-GPU sum in tests/test_dp.py -- 256 against
    128 FP32 accumulator updates of possibly truncating tensor-core adds stay below ~2e-5);
  * the same step on two nets built alike is bit-identical (fixed tile order, fixed reduction order).
The data-parallel tests (tests/test_dp.py) check the same sum across GPUs at small sizes."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DIMS, B = [4096, 4096, 4096, 1000], 4096


def make(g, precision):
    return g.SoftmaxCrossEntropyNeuralNet(DIMS, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=precision,
                                          flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_c4_full_batch_update_is_the_mean_of_its_halves(g, prec):
    precision = g.capi.PREC_BF16 if prec == "bf16" else g.capi.PREC_TF32
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
