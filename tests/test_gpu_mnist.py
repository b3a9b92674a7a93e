"""End-to-end MNISTTrainer path on the GPU (FP64 parity mode) against the oracle run on the same
data in the same order: IDX parse -> /255.0 -> train with the Random(1) sampler -> `>=` argmax
accuracy over both partitions -> trainLog.csv row.  The image files are the deterministic
MNIST-shaped SURROGATE (the reference mount lacks the real image files); the labels are the
reference's real label files (tests/golden/mnist)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def data_dir(tmp_path_factory, golden_dir):
    from gnn_b200 import mnist
    d = tmp_path_factory.mktemp("mnist_data")
    # small slices of the real label files keep the CPU oracle leg short
    for lab, img, n, seed in (("train-labels-idx1-ubyte", "train-images-idx3-ubyte", 3000, 1),
                              ("t10k-labels-idx1-ubyte", "t10k-images-idx3-ubyte", 1000, 2)):
        labels = mnist.read_idx_labels(os.path.join(golden_dir, "mnist", lab))[:n]
        import struct
        with open(d / lab, "wb") as f:
            f.write(struct.pack(">ii", 2049, n))
            f.write(labels.astype(np.uint8).tobytes())
        mnist.write_surrogate_images(labels, str(d / img), seed)
    return str(d)


@pytest.mark.parametrize("dims,B,step", [([784, 100, 50, 10], 8, 0.01), ([784, 300, 100, 10], 16, 0.0125)])
def test_mnist_trainer_matches_oracle(g, orc, data_dir, tmp_path, dims, B, step):
    from gnn_b200 import mnist
    iters = 400
    net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)   # reference init, undamped
    tr = g.MNISTTrainer(net, data_dir, str(tmp_path / "logs"))
    tr.train(iters, step, B, 0.9, False, False, False)
    train_acc, test_acc = tr.testOnTrainingData(), tr.testOnTestData()
    tr.logTest(dims, iters, step, B, 0.9, False, train_acc, test_acc)

    xtr = orc.read_idx_images(os.path.join(data_dir, "train-images-idx3-ubyte"))
    ltr = orc.read_idx_labels(os.path.join(data_dir, "train-labels-idx1-ubyte"))
    xte = orc.read_idx_images(os.path.join(data_dir, "t10k-images-idx3-ubyte"))
    lte = orc.read_idx_labels(os.path.join(data_dir, "t10k-labels-idx1-ubyte"))
    ref = orc.OracleNet(dims)
    ref.train(xtr, mnist.one_hot(ltr), iters, B, step, 0.9)
    rh_tr, rp_tr = ref.evaluate(xtr, ltr)
    rh_te, rp_te = ref.evaluate(xte, lte)

    wr = ref.get_weights()
    assert np.max(np.abs(net.getWeights() - wr) / np.maximum(1e-3, np.abs(wr))) <= 1e-9
    _, _, p_te = tr.testingPartition.evaluate(want_pred=True)
    _, _, p_tr = tr.trainer.dataMaster.evaluate(want_pred=True)
    assert (p_te == rp_te).all() and (p_tr == rp_tr).all()          # bit-exact argmax predictions
    assert train_acc == rh_tr / len(ltr) and test_acc == rh_te / len(lte)
    assert test_acc > 0.5                                            # it actually learned the surrogate
    row = open(tmp_path / "logs" / "trainLog.csv").read()
    assert row == orc.format_log_row(dims, iters, step, B, 0.9, False, rh_tr / len(ltr), rh_te / len(lte))


@pytest.fixture(scope="module")
def full_data_dir(tmp_path_factory, golden_dir):
    """all 60 000 + 10 000 rows: the reference's real label files with the surrogate images"""
    from gnn_b200 import mnist
    d = tmp_path_factory.mktemp("mnist_full")
    mnist.ensure_surrogate(str(d), os.path.join(golden_dir, "mnist"))
    return str(d)


def test_c2_one_full_epoch_matches_oracle(g, orc, full_data_dir, tmp_path):
    """BASELINE configs[1] across an epoch boundary: 60000 / 16 = 3750 gradient steps of 784-300-100-10 with the literals
    of MNISTTrainer.main (M:224-230) -- the sampler refills once the epoch is used up (T:147-150) -- plus ten steps of the
    second epoch, GPU (K1) against the oracle on the same rows, each side drawing its batches with its OWN sampler.

    Training from the reference's undamped U[-0.5,0.5) init is chaotic: a 1e-13 perturbation of the initial weights of
    the ORACLE ITSELF grows to 2e-9 (Frobenius) / 4e-6 (max relative) by step 2500 and ~1e-5 / O(1) by the end of the
    epoch (profiles/r02_notes.md), so no two implementations whose dot products round differently can agree to 1e-9 after
    a free-running epoch.  The epoch is therefore checked in six segments (5 x 750 steps + the 10 steps that cross the
    epoch boundary): after each segment the weights agree within 1e-9 relative, then the oracle adopts the GPU's weights
    and velocity and both go on -- every step of the epoch sits inside a tight comparison.  At the end: validate(601)
    (T:65, T:102-113) within 1e-9, argmax identical on all 10 000 test and 60 000 training rows, trainLog.csv row
    byte-identical.  SURROGATE images, not MNIST."""
    from gnn_b200 import mnist
    dims, B, step = [784, 300, 100, 10], 16, 0.0125
    segments = [750] * 5 + [10]
    iters = sum(segments)
    net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    tr = g.MNISTTrainer(net, full_data_dir, str(tmp_path / "logs"))
    assert tr.n_train == 60000 and tr.n_test == 10000
    xtr = orc.read_idx_images(os.path.join(full_data_dir, "train-images-idx3-ubyte"))
    ltr = orc.read_idx_labels(os.path.join(full_data_dir, "train-labels-idx1-ubyte"))
    xte = orc.read_idx_images(os.path.join(full_data_dir, "t10k-images-idx3-ubyte"))
    lte = orc.read_idx_labels(os.path.join(full_data_dir, "t10k-labels-idx1-ubyte"))
    ttr = mnist.one_hot(ltr)
    ref = orc.OracleNet(dims)
    ref_sampler = orc.Sampler(60000, seed=1)
    orc.set_threads(os.cpu_count() or 1)
    try:
        worst = 0.0
        for n in segments:
            tr.train(n, step, B, 0.9, False, False, False)                      # the GPU trainer's own sample() stream
            idx = np.stack([ref_sampler.sample(B) for _ in range(n)]).astype(np.int32)
            ref.train(xtr, ttr, n, B, step, 0.9, idx=idx)
            w, wr = net.getWeights(), ref.get_weights()
            rel = np.max(np.abs(w - wr) / np.maximum(1e-3, np.abs(wr)))
            worst = max(worst, rel)
            assert rel <= 1e-9, (n, rel)
            ref.set_weights(w)
            ref.set_velocity(net.getVelocity())
        loss_gpu = tr.trainer.validate(601)
        loss_ref = ref.validate(xtr, ttr, 601)
        rh_tr, rp_tr = ref.evaluate(xtr, ltr)
        rh_te, rp_te = ref.evaluate(xte, lte)
    finally:
        orc.set_threads(1)
    print("worst segment deviation %.3g" % worst)
    train_acc, test_acc = tr.testOnTrainingData(), tr.testOnTestData()
    tr.logTest(dims, iters, step, B, 0.9, False, train_acc, test_acc)
    assert abs(loss_gpu - loss_ref) <= 1e-9 * max(1.0, abs(loss_ref)), (loss_gpu, loss_ref)
    _, _, p_te = tr.testingPartition.evaluate(want_pred=True)
    _, _, p_tr = tr.trainer.dataMaster.evaluate(want_pred=True)
    assert (p_te == rp_te).all() and (p_tr == rp_tr).all()
    assert train_acc == rh_tr / 60000 and test_acc == rh_te / 10000
    row = open(tmp_path / "logs" / "trainLog.csv").read()
    assert row == orc.format_log_row(dims, iters, step, B, 0.9, False, rh_tr / 60000, rh_te / 10000)
