"""Host-side mirror of `NeuralNetTrainer` (src/NeuralNetTrainer.java:11-170) and `ProgressBar`
(src/ProgressBar.java) over the C-ABI fast path: the data set is staged once on the GPU
(dataMaster, T:35-40), sample() keeps the reference's draw semantics (T:143-168, native Fenwick-tree
sampler) and `train` ships whole runs of index lists per native call instead of one JNI crossing
per gradient step."""
import ctypes as C
import sys
import time

import numpy as np

from . import _capi as capi

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class ProgressBar:
    """Console ticker in the spirit of src/ProgressBar.java:117-139 (console UI only)."""

    def __init__(self, ticks, total, out=sys.stdout):
        self.ticks, self.total, self.out = ticks, total, out
        self.done, self.printed = 0, 0
        self.start = time.time()
        self.out.write("Starting...\n")

    def step(self, n=1):
        self.done += n
        while self.printed < self.ticks and self.done * self.ticks >= (self.printed + 1) * self.total:
            self.out.write("▯")
            self.out.flush()
            self.printed += 1

    def finish(self):
        self.out.write("\nFinished in %.2f seconds\n" % (time.time() - self.start))


class Dataset:
    """Rows + targets staged for a net (b2nn_dataset)."""

    def __init__(self, net, x, t=None, labels=None, location=capi.DATA_DEVICE):
        x = np.ascontiguousarray(x)
        if x.dtype == np.float64:
            dt = capi.DT_F64
        elif x.dtype == np.float32:
            dt = capi.DT_F32
        elif x.dtype == np.uint16:
            dt = capi.DT_BF16
        elif x.dtype == np.uint8:
            dt = capi.DT_U8        # IDX pixel bytes, read as (raw & 0xff) / 255.0 (M:97)
        else:
            raise ValueError("x must be float64, float32, uint16 (bfloat16 bits) or uint8 (pixel bytes)")
        if x.ndim != 2 or x.shape[1] != net.layerDims[0]:
            raise ValueError("data rows must have the net's input dimension (T:32)")
        tp = lp = None
        if t is not None:
            t = np.ascontiguousarray(t, dtype=np.float64)
            if t.shape != (x.shape[0], net.layerDims[-1]):
                raise ValueError("expected rows must have the net's output dimension (T:33)")
            tp = t.ctypes.data_as(_dp)
        if labels is not None:
            labels = np.ascontiguousarray(labels, dtype=np.int32)
            if labels.shape != (x.shape[0],):
                raise ValueError("one label per row")
            lp = labels.ctypes.data_as(_ip)
        h = C.c_void_p()
        capi.check(capi.lib().b2nn_dataset_create(net._h, x.ctypes.data_as(C.c_void_p), dt, tp, lp, x.shape[0],
                                                  location, C.byref(h)))
        self._h, self.net, self.n = h, net, x.shape[0]

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().b2nn_dataset_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def evaluate(self, n=None, want_loss=True, want_pred=False):
        n = self.n if n is None else n
        hits, loss = C.c_int64(0), C.c_double(0.0)
        pred = np.empty(n, dtype=np.int32) if want_pred else None
        capi.check(capi.lib().b2nn_evaluate(self.net._h, self._h, n, C.byref(hits),
                                            C.byref(loss) if want_loss else None,
                                            pred.ctypes.data_as(_ip) if want_pred else None))
        return hits.value, (loss.value if want_loss else None), pred


class Sampler:
    """sample()/refillSampler() (T:143-168) with java.util.Random(seed) draws."""

    def __init__(self, n_rows, seed=1):
        h = C.c_void_p()
        capi.check(capi.lib().b2nn_sampler_create(n_rows, seed, C.byref(h)))
        self._h = h

    def draw(self, iters, B):
        out = np.empty((iters, B), dtype=np.int32)
        capi.check(capi.lib().b2nn_sampler_draw(self._h, iters, B, out.ctypes.data_as(_ip)))
        return out

    def __del__(self):
        if getattr(self, "_h", None):
            capi.lib().b2nn_sampler_destroy(self._h)
            self._h = None


class NeuralNetTrainer:
    """NeuralNetTrainer(data, net) (T:28).  `data` is the Map<double[],double[]> given as an (X, T)
    pair of 2-D arrays or a sequence of (input, expected) pairs; the master order (the reference's
    HashMap order is identity-hash dependent) is the row order."""

    CHUNK = 2000  # iterations per native call when only a progress monitor is attached

    def __init__(self, data, net):
        if data is None or net is None:
            raise ValueError("data and net must be non-null (T:30-31)")
        if isinstance(data, tuple) and len(data) == 2 and np.ndim(data[0]) == 2:
            x, t = data
        else:
            pairs = list(data.items()) if hasattr(data, "items") else list(data)
            x = [p[0] for p in pairs]
            t = [p[1] for p in pairs]
        x = np.ascontiguousarray(x, dtype=np.float64)
        t = np.ascontiguousarray(t, dtype=np.float64)
        if x.shape[1] != net.getInputDim() or t.shape[1] != net.getOutputDim():
            raise ValueError("data dimensions must match the net (T:32-33)")
        self.net = net
        self.dataMaster = Dataset(net, x, t=t)         # T:35-39
        self.sampler = Sampler(x.shape[0], seed=1)      # refillSampler T:40, Random(1) T:42
        self.lastIndices = None

    def size(self):
        return self.dataMaster.n

    def train(self, iterations, stepSize, batchSize, momentum, noise=False, monitor=None, observer=None):
        """T:60-92."""
        if not (iterations > 0 and stepSize > 0):
            raise ValueError("iterations and stepSize must be positive (T:62)")
        if not (0 < batchSize < self.dataMaster.n):
            raise ValueError("batchSize must be positive and below the data size (T:63)")
        if noise:
            raise NotImplementedError("noise=true is not supported by the GPU backend (NaN-producing term, S:335)")
        validationSize = self.dataMaster.n // 100 + 1     # T:65
        L = capi.lib()
        if observer is not None:
            for i in range(iterations):                   # T:69-73 / T:76-79
                idx = self.sampler.draw(1, batchSize)
                capi.check(L.b2nn_train_indexed(self.net._h, self.dataMaster._h, idx.ctypes.data_as(_ip), 1,
                                                batchSize, stepSize, momentum, None))
                observer.write("%d,%.2f\n" % (i, self.validate(validationSize)))
                if monitor is not None:
                    monitor.step()
        else:
            done = 0
            while done < iterations:                      # T:82-86 / T:88-90
                n = min(self.CHUNK if monitor is not None else 20000, iterations - done)
                idx = self.sampler.draw(n, batchSize)
                capi.check(L.b2nn_train_indexed(self.net._h, self.dataMaster._h, idx.ctypes.data_as(_ip), n,
                                                batchSize, stepSize, momentum, None))
                done += n
                if monitor is not None:
                    self.net.sync()
                    monitor.step(n)
        self.net.sync()
        if monitor is not None:
            monitor.finish()

    def trainIndexed(self, idx, stepSize, momentum, wantLoss=False):
        """the fast path with caller-drawn batches: idx [iters][B]."""
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        iters, B = idx.shape
        loss = np.empty(iters, dtype=np.float64) if wantLoss else None
        capi.check(capi.lib().b2nn_train_indexed(self.net._h, self.dataMaster._h, idx.ctypes.data_as(_ip), iters, B,
                                                 stepSize, momentum, loss.ctypes.data_as(_dp) if wantLoss else None))
        return loss

    def validate(self, batchSize):
        """validate(int), T:102-113: mean loss over the first batchSize rows of dataMaster."""
        if not (0 < batchSize < self.dataMaster.n):
            raise ValueError("batchSize must be positive and below the data size (T:104)")
        out = C.c_double(0.0)
        capi.check(capi.lib().b2nn_validate(self.net._h, self.dataMaster._h, batchSize, C.byref(out)))
        return out.value
