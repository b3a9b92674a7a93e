"""Host-side mirror of `MNISTTrainer` (src/MNISTTrainer.java:11-246): IDX parsing with the /255.0
normalisation (M:76-118), training through NeuralNetTrainer, `>=`-argmax accuracy over the training
and test partitions (M:159-197) evaluated on the GPU, and the logs/trainLog.csv row (M:211-219).
Also a deterministic MNIST-shaped surrogate image generator: the reference mount lacks the two
image files (.MISSING_LARGE_BLOBS), only the label files are real."""
import os
import struct

import numpy as np

from . import _capi as capi
from . import functions as F
from .nets import SoftmaxCrossEntropyNeuralNet
from .trainer import Dataset, NeuralNetTrainer, ProgressBar


def read_idx_images(path):
    """M:39-45,92-100: big-endian header (magic 2051 checked unconditionally),
    pixel (raw & 0xff) / 255.0."""
    with open(path, "rb") as f:
        magic, n, h, w = struct.unpack(">iiii", f.read(16))
        if magic != 2051:
            raise IOError("%s: bad IDX3 magic %d" % (path, magic))
        raw = np.frombuffer(f.read(n * h * w), dtype=np.uint8)
    if raw.size != n * h * w:
        raise IOError("%s: truncated" % path)
    return raw.reshape(n, h * w).astype(np.float64) / 255.0


def read_idx_labels(path):
    """M:38,40,111-118."""
    with open(path, "rb") as f:
        magic, n = struct.unpack(">ii", f.read(8))
        if magic != 2049:
            raise IOError("%s: bad IDX1 magic %d" % (path, magic))
        raw = np.frombuffer(f.read(n), dtype=np.uint8)
    if raw.size != n:
        raise IOError("%s: truncated" % path)
    return raw.astype(np.int32)


def one_hot(labels, n_classes=10):
    """readLabel, M:111-118."""
    t = np.zeros((labels.shape[0], n_classes), dtype=np.float64)
    t[np.arange(labels.shape[0]), labels] = 1.0
    return t


def write_surrogate_images(labels, path, seed):
    """Deterministic MNIST-shaped stand-in images in IDX3 format: a smooth per-class prototype,
    shifted by up to +-2 pixels and perturbed by seeded noise, quantised to uint8.  NOT MNIST."""
    rs = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:28, 0:28]
    protos = np.zeros((10, 28, 28))
    prs = np.random.RandomState(1234)  # prototypes are shared by the train and test files
    for c in range(10):
        for _ in range(4):
            cy, cx = prs.uniform(6, 22, size=2)
            sy, sx = prs.uniform(1.5, 4.0, size=2)
            protos[c] += np.exp(-((yy - cy) ** 2 / (2 * sy * sy) + (xx - cx) ** 2 / (2 * sx * sx)))
        protos[c] /= protos[c].max()
    n = labels.shape[0]
    out = np.empty((n, 28, 28), dtype=np.uint8)
    dy = rs.randint(-2, 3, size=n)
    dx = rs.randint(-2, 3, size=n)
    for i in range(n):
        img = np.roll(np.roll(protos[labels[i]], dy[i], axis=0), dx[i], axis=1)
        img = img * rs.uniform(0.6, 1.0) + rs.normal(0.0, 0.25, size=(28, 28))
        out[i] = np.clip(img * 255.0, 0, 255).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(struct.pack(">iiii", 2051, n, 28, 28))
        f.write(out.tobytes())


def ensure_surrogate(data_dir, label_dir):
    """Place label files + surrogate image files under data_dir (idempotent)."""
    os.makedirs(data_dir, exist_ok=True)
    for lab, img, seed in (("train-labels-idx1-ubyte", "train-images-idx3-ubyte", 1),
                           ("t10k-labels-idx1-ubyte", "t10k-images-idx3-ubyte", 2)):
        lp = os.path.join(data_dir, lab)
        if not os.path.exists(lp):
            with open(os.path.join(label_dir, lab), "rb") as src, open(lp, "wb") as dst:
                dst.write(src.read())
        ip = os.path.join(data_dir, img)
        if not os.path.exists(ip):
            write_surrogate_images(read_idx_labels(lp), ip, seed)


class MNISTTrainer:
    """MNISTTrainer(net) (M:26): data/ and logs/ are CWD-relative in the reference (M:28-31,61)."""

    def __init__(self, net, data_dir="data", log_dir="logs", max_train=None, max_test=None):
        if net is None:
            raise ValueError("net must be non-null (M:63)")
        trl = read_idx_labels(os.path.join(data_dir, "train-labels-idx1-ubyte"))
        tri = read_idx_images(os.path.join(data_dir, "train-images-idx3-ubyte"))
        tel = read_idx_labels(os.path.join(data_dir, "t10k-labels-idx1-ubyte"))
        tei = read_idx_images(os.path.join(data_dir, "t10k-images-idx3-ubyte"))
        if trl.shape[0] != tri.shape[0] or tel.shape[0] != tei.shape[0]:
            raise IOError("label / image counts differ (M:42,54)")
        if tri.shape[1] != tei.shape[1]:
            raise IOError("training and test image sizes differ (M:56)")
        if max_train:
            trl, tri = trl[:max_train], tri[:max_train]
        if max_test:
            tel, tei = tel[:max_test], tei[:max_test]
        self.trainLabels, self.testLabels = trl, tel
        os.makedirs(log_dir, exist_ok=True)
        self.trainingLog = open(os.path.join(log_dir, "trainLog.csv"), "a")   # append mode, M:61 (Q18)
        self.log_dir = log_dir
        self.net = net
        self.trainer = NeuralNetTrainer((tri, one_hot(trl)), net)             # M:65
        self.testingPartition = Dataset(net, tei, labels=tel)
        self.n_train, self.n_test = tri.shape[0], tei.shape[0]

    def train(self, iterations, stepSize, batchSize, momentum, noise, monitoring, logging):
        """M:139-151"""
        pb = ProgressBar(10, iterations) if monitoring else None
        log = None
        if logging:
            import time
            ident = time.time_ns() % 99999                                    # M:146 (Q19)
            log = open(os.path.join(self.log_dir, "log%d.csv" % ident), "w")
            print("Logging net performance to %d" % ident)
        try:
            self.trainer.train(iterations, stepSize, batchSize, momentum, noise, pb, log)
        finally:
            if log:
                log.close()

    def testOnTestData(self):
        """M:159-172: hits / size with the `>=` argmax."""
        hits, _, _ = self.testingPartition.evaluate(want_loss=False)
        return hits / self.n_test

    def testOnTrainingData(self):
        """M:180-197"""
        hits, _, _ = self.trainer.dataMaster.evaluate(want_loss=False)
        return hits / self.n_train

    @staticmethod
    def formatLogRow(netDim, iterations, stepSize, batchSize, momentum, noise, trainingAcc, testAcc):
        """M:211-219"""
        return "%s,%d,%.5f,%d,%.3f,%s,%.5f,%.5f\n" % ("-".join(str(d) for d in netDim), iterations, stepSize,
                                                     batchSize, momentum, "true" if noise else "false",
                                                     trainingAcc, testAcc)

    def logTest(self, netDim, iterations, stepSize, batchSize, momentum, noise, trainingAcc, testAcc):
        self.trainingLog.write(self.formatLogRow(netDim, iterations, stepSize, batchSize, momentum, noise,
                                                 trainingAcc, testAcc))
        self.trainingLog.flush()


def main(argv=None, data_dir="data", log_dir="logs", iterations=100000):
    """MNISTTrainer.main (M:221-245) with its literals; the two inline lambdas become the
    LEAKY_RELU singletons (the one edit the drop-in needs)."""
    netDim = [784, 300, 100, 10]
    stepSize, batchSize, momentum, noise = 0.0125, 16, 0.9, False
    net = SoftmaxCrossEntropyNeuralNet(netDim, F.LEAKY_RELU, F.LEAKY_RELU_PRIME)
    trainer = MNISTTrainer(net, data_dir, log_dir)
    trainer.train(iterations, stepSize, batchSize, momentum, noise, True, False)
    trainAcc = trainer.testOnTrainingData()     # evaluation order: train first, then test (M:241)
    testAcc = trainer.testOnTestData()
    trainer.logTest(netDim, iterations, stepSize, batchSize, momentum, noise, trainAcc, testAcc)
    return trainAcc, testAcc
