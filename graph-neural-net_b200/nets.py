"""Host-side mirror of the reference's net classes over the C-ABI: `NeuralNet`
(src/NeuralNet.java:7-67), `SoftmaxCrossEntropyNeuralNet` (src/SoftmaxCrossEntropyNeuralNet.java:11)
and `GeneralNeuralNet` (src/GeneralNeuralNet.java:11).  Same method names, argument meaning and
error behaviour; every body is one call into libb2nn.so."""
import ctypes as C

import numpy as np

from . import _capi as capi
from . import functions as F

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _d(a):
    return a.ctypes.data_as(_dp)


class NeuralNet:
    """The 6-method contract (src/NeuralNet.java:16-65) plus the native handle."""

    def __init__(self, layerDims, kind, inner, last, loss, precision, flags, seed, device):
        if layerDims is None or len(layerDims) < 2:
            raise ValueError("layerDims must be non-null with at least 2 entries (S:105)")
        self.layerDims = [int(d) for d in layerDims]
        dims = np.asarray(self.layerDims, dtype=np.int32)
        h = C.c_void_p()
        capi.check(capi.lib().b2nn_create(dims.ctypes.data_as(_ip), len(self.layerDims), kind, inner[0], inner[1],
                                          last[0], last[1], loss, precision, flags, seed, device, C.byref(h)))
        self._h = h
        self.precision = precision
        n = C.c_int64()
        capi.check(capi.lib().b2nn_num_weights(self._h, C.byref(n)))
        self.numWeights = n.value

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().b2nn_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- NeuralNet interface -------------------------------------------------------------
    def propagate(self, input):
        """src/NeuralNet.java:16; also accepts [n][d0] and returns [n][dL]."""
        x = _f64(input)
        single = x.ndim == 1
        x = x.reshape(-1, x.shape[-1])
        if x.shape[1] != self.layerDims[0]:
            raise ValueError("input length %d != input dimension %d (S:166)" % (x.shape[1], self.layerDims[0]))
        y = np.empty((x.shape[0], self.layerDims[-1]), dtype=np.float64)
        capi.check(capi.lib().b2nn_propagate(self._h, _d(x), x.shape[0], _d(y)))
        return y[0] if single else y

    def propagateDeviceNoAlloc(self, x_dev_ptr, n, y_dev_ptr):
        """b2nn_propagate_device_noalloc (no reference counterpart): propagate for FP64 nets over DEVICE buffers
        (addresses as ints: x [n][d0] in, y [n][dL] out), work rows in the library's fixed 4 MB device array, nothing
        allocated, launched on the net's stream without a sync (call sync())."""
        capi.check(capi.lib().b2nn_propagate_device_noalloc(self._h, C.c_void_p(int(x_dev_ptr)), int(n),
                                                            C.c_void_p(int(y_dev_ptr))))

    def calculateLoss(self, input, expected):
        """src/NeuralNet.java:27; batched inputs return one loss per row."""
        x, t = _f64(input), _f64(expected)
        single = x.ndim == 1
        x = x.reshape(-1, x.shape[-1])
        t = t.reshape(-1, t.shape[-1])
        if x.shape[1] != self.layerDims[0] or t.shape[1] != self.layerDims[-1] or x.shape[0] != t.shape[0]:
            raise ValueError("input/expected dimensions do not match the net (S:210)")
        out = np.empty(x.shape[0], dtype=np.float64)
        capi.check(capi.lib().b2nn_loss(self._h, _d(x), _d(t), x.shape[0], _d(out)))
        return float(out[0]) if single else out

    def calculateWeightGradient(self, input, expected):
        """src/NeuralNet.java:39: {layer i: [in][out] matrix}."""
        x, t = _f64(input), _f64(expected)
        if x.shape != (self.layerDims[0],) or t.shape != (self.layerDims[-1],):
            raise ValueError("input/expected dimensions do not match the net (S:232)")
        g = np.empty(self.numWeights, dtype=np.float64)
        capi.check(capi.lib().b2nn_weight_gradient(self._h, _d(x), _d(t), _d(g)))
        out, off = {}, 0
        for l in range(len(self.layerDims) - 1):
            n = self.layerDims[l] * self.layerDims[l + 1]
            out[l] = g[off:off + n].reshape(self.layerDims[l], self.layerDims[l + 1])
            off += n
        return out

    def gradientStep(self, batch, step, momentum, noise=False, returnLoss=False):
        """src/NeuralNet.java:51.  `batch` is the Map<double[],double[]> as a sequence of
        (input, expected) pairs or an (X, T) pair of 2-D arrays; iteration order = row order."""
        if batch is None:
            raise ValueError("batch must be non-null (S:299)")
        if isinstance(batch, tuple) and len(batch) == 2 and np.ndim(batch[0]) == 2:
            x, t = _f64(batch[0]), _f64(batch[1])
        else:
            pairs = list(batch.items()) if hasattr(batch, "items") else list(batch)
            if not pairs:
                raise ValueError("batch must not be empty (S:300)")
            x = _f64([p[0] for p in pairs])
            t = _f64([p[1] for p in pairs])
        if x.shape[0] == 0:
            raise ValueError("batch must not be empty (S:300)")
        if x.shape[1] != self.layerDims[0] or t.shape[1] != self.layerDims[-1] or x.shape[0] != t.shape[0]:
            raise ValueError("batch dimensions do not match the net")
        loss = C.c_double(0.0)
        capi.check(capi.lib().b2nn_gradient_step(self._h, _d(x), _d(t), x.shape[0], step, momentum, int(bool(noise)),
                                                 C.byref(loss) if returnLoss else None))
        return loss.value if returnLoss else None

    def getInputDim(self):
        i = C.c_int32()
        capi.check(capi.lib().b2nn_dims(self._h, C.byref(i), None))
        return i.value

    def getOutputDim(self):
        o = C.c_int32()
        capi.check(capi.lib().b2nn_dims(self._h, None, C.byref(o)))
        return o.value

    # ---- hooks the reference lacks (checkpoint / parity tests) ---------------------------
    def getWeights(self):
        w = np.empty(self.numWeights, dtype=np.float64)
        capi.check(capi.lib().b2nn_get_weights(self._h, _d(w)))
        return w

    def setWeights(self, w):
        w = _f64(w)
        if w.size != self.numWeights:
            raise ValueError("expected %d weights" % self.numWeights)
        capi.check(capi.lib().b2nn_set_weights(self._h, _d(w)))

    def getVelocity(self):
        v = np.empty(self.numWeights, dtype=np.float64)
        capi.check(capi.lib().b2nn_get_velocity(self._h, _d(v)))
        return v

    def setVelocity(self, v):
        v = _f64(v)
        if v.size != self.numWeights:
            raise ValueError("expected %d entries" % self.numWeights)
        capi.check(capi.lib().b2nn_set_velocity(self._h, _d(v)))

    def getTime(self):
        """`time` (S:23): gradient steps taken (S:343)."""
        t = C.c_int64(0)
        capi.check(capi.lib().b2nn_time(self._h, C.byref(t)))
        return t.value

    def saveCheckpoint(self, path):
        """weights, previousUpdate and time to a flat file (csrc/checkpoint.cu); the reference has no persistence."""
        import os
        capi.check(capi.lib().b2nn_checkpoint_save(self._h, os.fsencode(path)))

    def loadCheckpoint(self, path):
        """restore what saveCheckpoint wrote; layerDims, class and functions must match (ValueError otherwise)."""
        import os
        capi.check(capi.lib().b2nn_checkpoint_load(self._h, os.fsencode(path)))

    def gradientStepLabeled(self, x, labels, step, momentum, returnLoss=False):
        """one-hot targets + narrow host dtypes (wide BF16 nets)."""
        x = np.ascontiguousarray(x)
        if x.dtype == np.float64:
            dt = capi.DT_F64
        elif x.dtype == np.float32:
            dt = capi.DT_F32
        elif x.dtype == np.uint16:  # raw bfloat16 bits
            dt = capi.DT_BF16
        else:
            raise ValueError("x must be float64, float32 or uint16 (bfloat16 bits)")
        labels = np.ascontiguousarray(labels, dtype=np.int32)
        if x.ndim != 2 or x.shape[1] != self.layerDims[0] or labels.shape != (x.shape[0],):
            raise ValueError("batch dimensions do not match the net")
        loss = C.c_double(0.0)
        capi.check(capi.lib().b2nn_gradient_step_labeled(self._h, x.ctypes.data_as(C.c_void_p), dt,
                                                         labels.ctypes.data_as(_ip), x.shape[0], step, momentum,
                                                         C.byref(loss) if returnLoss else None))
        return loss.value if returnLoss else None

    def sync(self):
        capi.check(capi.lib().b2nn_sync(self._h))

    def setStream(self, cuda_stream_ptr):
        capi.check(capi.lib().b2nn_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def kernelLaunches(self):
        n = C.c_int64()
        capi.check(capi.lib().b2nn_kernel_launches(self._h, C.byref(n)))
        return n.value

    def profileEnable(self, on=True):
        capi.check(capi.lib().b2nn_profile_enable(self._h, int(on)))

    def profileRead(self):
        """(gemm_ms, gemm_launches, gemm_flops) since profiling was enabled"""
        ms, n, fl = C.c_double(), C.c_int64(), C.c_double()
        capi.check(capi.lib().b2nn_profile_read(self._h, C.byref(ms), C.byref(n), C.byref(fl)))
        return ms.value, n.value, fl.value

    K1_PHASES = ("stage", "phase1", "barrier1", "phase2", "barrier2", "p3:delta2", "phase4", "p5:W2", "p3:a2", "p3:z3",
                 "p3:softmax", "p3:loss", "p5:W0", "p5:W1c", "p5:W1r", "stage:wait")

    def k1Profile(self, on=True):
        """switch K1's per-phase cycle counters on / off; returns ({phase: cycles per step}, steps) of the last
        profiled launch"""
        cyc = (C.c_double * 16)()
        n = C.c_int32(0)
        capi.check(capi.lib().b2nn_k1_profile(self._h, int(on), cyc, C.byref(n)))
        return dict(zip(self.K1_PHASES, [float(c) for c in cyc])), n.value

    def profileReadStamps(self):
        """(busy_ms, sum_ms, launches, flops) of the contraction launches since profileEnable(3): in-kernel %globaltimer
        stamps, no events, nothing serialised"""
        busy, tot, n, fl = C.c_double(), C.c_double(), C.c_int64(), C.c_double()
        capi.check(capi.lib().b2nn_profile_read_stamps(self._h, C.byref(busy), C.byref(tot), C.byref(n), C.byref(fl)))
        return busy.value, tot.value, n.value, fl.value

    def profileStampClockMHz(self):
        """SM clock (MHz) the launches of the last profileReadStamps() ran at: cycles / ns of CTA 0 of every launch"""
        mhz = C.c_double()
        capi.check(capi.lib().b2nn_profile_stamp_clock_mhz(self._h, C.byref(mhz)))
        return mhz.value

    def profileReadClasses(self):
        """{class: (ms, launches)} of the profiled pass: contraction, update, output, prologue"""
        ms, n = (C.c_double * 4)(), (C.c_int64 * 4)()
        capi.check(capi.lib().b2nn_profile_read_classes(self._h, ms, n))
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(("contraction", "update", "output", "prologue"))}

    def commInit(self, unique_id, rank, world):
        buf = C.create_string_buffer(bytes(unique_id), 128)
        capi.check(capi.lib().b2nn_comm_init(self._h, buf, rank, world))


    def commIpcExport(self):
        buf = C.create_string_buffer(64)
        capi.check(capi.lib().b2nn_comm_ipc_export(self._h, buf))
        return buf.raw

    def commIpcImport(self, handles):
        blob = b"".join(bytes(h) for h in handles)
        buf = C.create_string_buffer(blob, len(blob))
        capi.check(capi.lib().b2nn_comm_ipc_import(self._h, buf, len(handles)))

    def commNvlsExport(self):
        """-> (pid, fd) of this rank's arena, or None when the device / driver has no multicast support"""
        pf = (C.c_int32 * 2)()
        rc = capi.lib().b2nn_comm_nvls_export(self._h, pf)
        if rc == capi.ERR_UNSUPPORTED:
            return None
        capi.check(rc)
        return int(pf[0]), int(pf[1])

    def commNvlsImport(self, pid_fds):
        flat = (C.c_int32 * (2 * len(pid_fds)))(*[int(x) for pf in pid_fds for x in pf])
        capi.check(capi.lib().b2nn_comm_nvls_import(self._h, flat, len(pid_fds)))

    def commNvlsTeam(self, pid_fd=None):
        """rank 0: pid_fd None -> returns the multicast object's (pid, fd); others pass rank 0's pair"""
        pf = (C.c_int32 * 2)(*(pid_fd or (0, 0)))
        capi.check(capi.lib().b2nn_comm_nvls_team(self._h, pf))
        return int(pf[0]), int(pf[1])

    def commMode(self):
        """'single' | 'nccl' | 'p2p' | 'nvls' -- the gradient exchange the next step uses"""
        m = C.c_int32()
        capi.check(capi.lib().b2nn_comm_mode(self._h, C.byref(m)))
        return ("single", "nccl", "p2p", "nvls")[m.value]

    def commNvlsBind(self):
        capi.check(capi.lib().b2nn_comm_nvls_bind(self._h))


def commUniqueId():
    buf = C.create_string_buffer(128)
    capi.check(capi.lib().b2nn_comm_unique_id(buf))
    return buf.raw


class SoftmaxCrossEntropyNeuralNet(NeuralNet):
    """src/SoftmaxCrossEntropyNeuralNet.java:103: (layerDims, innerActivationFunc, innerActivationPrime)."""

    def __init__(self, layerDims, innerActivationFunc, innerActivationPrime, precision=capi.PREC_F64, flags=0,
                 seed=1, device=-1):
        if innerActivationFunc is None or innerActivationPrime is None:
            raise ValueError("activation functions must be non-null (S:106)")
        inner = F.activation_pair_id(innerActivationFunc, innerActivationPrime, "inner")
        super().__init__(layerDims, capi.NET_SOFTMAX_CE, inner, (0, 0.0), 0, precision, flags, seed, device)


class GeneralNeuralNet(NeuralNet):
    """src/GeneralNeuralNet.java:112-115: (layerDims, inner f/f', last f/f', loss, loss')."""

    def __init__(self, layerDims, innerActivationFunc, innerActivationPrime, lastActivationFunc, lastActivationPrime,
                 lossFunc, lossPrime, precision=capi.PREC_F64, flags=0, seed=1, device=-1):
        for f in (innerActivationFunc, innerActivationPrime, lastActivationFunc, lastActivationPrime, lossFunc,
                  lossPrime):
            if f is None:
                raise ValueError("functions must be non-null (G:119-122)")
        inner = F.activation_pair_id(innerActivationFunc, innerActivationPrime, "inner")
        last = F.activation_pair_id(lastActivationFunc, lastActivationPrime, "last")
        loss = F.loss_pair_id(lossFunc, lossPrime)
        super().__init__(layerDims, capi.NET_GENERAL, inner, last, loss, precision, flags, seed, device)
