"""ctypes loader for the CPU oracle (TEST INFRASTRUCTURE, NOT PRODUCT CODE).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product package never does.
Parity status: see oracle/b2nn_oracle.h (JavaRandom pinned by KATs; the net
arithmetic is "parity unpinned" -- no JVM and no reference tests exist).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

ACT = {"leaky_relu": 0, "sigmoid": 1, "relu": 2, "tanh": 3, "identity": 4}
LOSS = {"half_squared_error": 0}
NET_SOFTMAX_CE, NET_GENERAL = 0, 1


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("b2nn_oracle.c", "b2nn_oracle_trainer.c", "b2nn_oracle.h")]
    if (not force and os.path.exists(_SO)
            and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs if os.path.exists(s))):
        return _SO
    subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _SO


class _JRandom(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("have", C.c_int), ("next_gaussian", C.c_double)]


_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(build())
    L.orc_jrandom_init.argtypes = [C.POINTER(_JRandom), C.c_int64]
    L.orc_jrandom_next_int.argtypes = [C.POINTER(_JRandom)]
    L.orc_jrandom_next_int.restype = C.c_int32
    L.orc_jrandom_next_int_bound.argtypes = [C.POINTER(_JRandom), C.c_int32]
    L.orc_jrandom_next_int_bound.restype = C.c_int32
    L.orc_jrandom_next_double.argtypes = [C.POINTER(_JRandom)]
    L.orc_jrandom_next_double.restype = C.c_double
    L.orc_jrandom_next_gaussian.argtypes = [C.POINTER(_JRandom)]
    L.orc_jrandom_next_gaussian.restype = C.c_double
    L.orc_jrandom_state.argtypes = [C.POINTER(_JRandom)]
    L.orc_jrandom_state.restype = C.c_uint64
    for name in ("orc_act", "orc_act_prime"):
        f = getattr(L, name)
        f.argtypes = [C.c_int, C.c_double, C.c_double]
        f.restype = C.c_double
    for name in ("orc_loss", "orc_loss_prime"):
        f = getattr(L, name)
        f.argtypes = [C.c_int, C.c_double, C.c_double]
        f.restype = C.c_double
    L.orc_net_create.argtypes = [_ip, C.c_int32, C.c_int32, C.c_int32, C.c_double,
                                 C.c_int32, C.c_double, C.c_int32, C.c_int64]
    L.orc_net_create.restype = C.c_void_p
    L.orc_net_destroy.argtypes = [C.c_void_p]
    L.orc_net_num_weights.argtypes = [C.c_void_p]
    L.orc_net_num_weights.restype = C.c_int64
    L.orc_net_input_dim.argtypes = [C.c_void_p]
    L.orc_net_output_dim.argtypes = [C.c_void_p]
    for name in ("orc_net_get_weights", "orc_net_set_weights", "orc_net_get_velocity", "orc_net_set_velocity"):
        getattr(L, name).argtypes = [C.c_void_p, _dp]
    L.orc_net_rng_state.argtypes = [C.c_void_p]
    L.orc_net_rng_state.restype = C.c_uint64
    L.orc_net_propagate.argtypes = [C.c_void_p, _dp, _dp]
    L.orc_net_loss.argtypes = [C.c_void_p, _dp, _dp]
    L.orc_net_loss.restype = C.c_double
    L.orc_net_weight_gradient.argtypes = [C.c_void_p, _dp, _dp, _dp]
    L.orc_net_gradient_step.argtypes = [C.c_void_p, _dp, _dp, C.c_int32, C.c_double, C.c_double,
                                        C.c_int32, C.c_int32]
    L.orc_set_threads.argtypes = [C.c_int32]
    L.orc_get_threads.restype = C.c_int32
    L.orc_sampler_create.argtypes = [C.c_int32, C.c_int64]
    L.orc_sampler_create.restype = C.c_void_p
    L.orc_sampler_destroy.argtypes = [C.c_void_p]
    L.orc_sampler_sample.argtypes = [C.c_void_p, C.c_int32, _ip]
    L.orc_train.argtypes = [C.c_void_p, _dp, _dp, C.c_int32, _ip, C.c_int32, C.c_int32,
                            C.c_double, C.c_double, C.c_int32, C.c_int32, _dp]
    L.orc_validate.argtypes = [C.c_void_p, _dp, _dp, C.c_int32]
    L.orc_validate.restype = C.c_double
    L.orc_evaluate.argtypes = [C.c_void_p, _dp, _ip, C.c_int64, _ip]
    L.orc_evaluate.restype = C.c_int64
    L.orc_read_idx_images.argtypes = [C.c_char_p, _dp, C.c_int64, _ip]
    L.orc_read_idx_images.restype = C.c_int64
    L.orc_read_idx_labels.argtypes = [C.c_char_p, _ip, C.c_int64]
    L.orc_read_idx_labels.restype = C.c_int64
    L.orc_format_log_row.argtypes = [C.c_char_p, C.c_int, _ip, C.c_int32, C.c_int32, C.c_double,
                                     C.c_int32, C.c_double, C.c_int32, C.c_double, C.c_double]
    _lib = L
    return L


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class JavaRandom:
    """java.util.Random restated (oracle/b2nn_oracle.c)."""

    def __init__(self, seed=1):
        self._s = _JRandom()
        lib().orc_jrandom_init(C.byref(self._s), seed)

    def next_int(self, bound=None):
        if bound is None:
            return lib().orc_jrandom_next_int(C.byref(self._s))
        return lib().orc_jrandom_next_int_bound(C.byref(self._s), bound)

    def next_double(self):
        return lib().orc_jrandom_next_double(C.byref(self._s))

    def next_gaussian(self):
        return lib().orc_jrandom_next_gaussian(C.byref(self._s))

    @property
    def state(self):
        return lib().orc_jrandom_state(C.byref(self._s))


class OracleNet:
    """SoftmaxCrossEntropyNeuralNet / GeneralNeuralNet restated on the CPU."""

    def __init__(self, dims, kind=NET_SOFTMAX_CE, inner_act="leaky_relu", inner_param=0.01,
                 last_act="sigmoid", last_param=0.0, loss="half_squared_error", seed=1):
        self.dims = [int(d) for d in dims]
        d = np.asarray(self.dims, dtype=np.int32)
        self._h = lib().orc_net_create(_i(d), len(self.dims), kind, ACT[inner_act], inner_param,
                                       ACT[last_act], last_param, LOSS[loss], seed)
        if not self._h:
            raise ValueError("orc_net_create failed")
        self.n_weights = lib().orc_net_num_weights(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_net_destroy(self._h)
            self._h = None

    def get_weights(self):
        out = np.empty(self.n_weights, dtype=np.float64)
        lib().orc_net_get_weights(self._h, _d(out))
        return out

    def set_weights(self, w):
        w = _f64(w)
        assert w.size == self.n_weights
        lib().orc_net_set_weights(self._h, _d(w))

    def get_velocity(self):
        out = np.empty(self.n_weights, dtype=np.float64)
        lib().orc_net_get_velocity(self._h, _d(out))
        return out

    def set_velocity(self, v):
        v = _f64(v)
        lib().orc_net_set_velocity(self._h, _d(v))

    @property
    def rng_state(self):
        return lib().orc_net_rng_state(self._h)

    def propagate(self, x):
        x = _f64(x)
        single = x.ndim == 1
        x2 = x.reshape(-1, self.dims[0])
        y = np.empty((x2.shape[0], self.dims[-1]), dtype=np.float64)
        for r in range(x2.shape[0]):
            xr = np.ascontiguousarray(x2[r])
            yr = np.empty(self.dims[-1], dtype=np.float64)
            lib().orc_net_propagate(self._h, _d(xr), _d(yr))
            y[r] = yr
        return y[0] if single else y

    def loss(self, x, t):
        x, t = _f64(x), _f64(t)
        return lib().orc_net_loss(self._h, _d(x), _d(t))

    def weight_gradient(self, x, t):
        x, t = _f64(x), _f64(t)
        g = np.empty(self.n_weights, dtype=np.float64)
        lib().orc_net_weight_gradient(self._h, _d(x), _d(t), _d(g))
        return g

    def gradient_step(self, x, t, step, momentum, noise=False, faithful=False):
        x, t = _f64(x), _f64(t)
        B = x.shape[0]
        lib().orc_net_gradient_step(self._h, _d(x), _d(t), B, step, momentum, int(noise), int(faithful))

    def train(self, x, t, iters, B, step, momentum, idx=None, noise=False, faithful=False, want_loss=False):
        x, t = _f64(x), _f64(t)
        losses = np.empty(iters, dtype=np.float64) if want_loss else None
        ip = None
        if idx is not None:
            idx = np.ascontiguousarray(idx, dtype=np.int32)
            assert idx.size == iters * B
            ip = _i(idx)
        lib().orc_train(self._h, _d(x), _d(t), x.shape[0], ip, iters, B, step, momentum,
                        int(noise), int(faithful), _d(losses) if want_loss else None)
        return losses

    def validate(self, x, t, count):
        x, t = _f64(x), _f64(t)
        return lib().orc_validate(self._h, _d(x), _d(t), count)

    def evaluate(self, x, labels):
        x = _f64(x)
        labels = np.ascontiguousarray(labels, dtype=np.int32)
        pred = np.empty(x.shape[0], dtype=np.int32)
        hits = lib().orc_evaluate(self._h, _d(x), _i(labels), x.shape[0], _i(pred))
        return hits, pred


def set_threads(t):
    """worker threads of the checker (bit-identical results; the reference and the timed CPU baseline are serial)"""
    lib().orc_set_threads(int(t))


def get_threads():
    return lib().orc_get_threads()


class Sampler:
    """NeuralNetTrainer.sample() restated; master order = row order."""

    def __init__(self, n_rows, seed=1):
        self._h = lib().orc_sampler_create(n_rows, seed)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_sampler_destroy(self._h)
            self._h = None

    def sample(self, B):
        out = np.empty(B, dtype=np.int32)
        lib().orc_sampler_sample(self._h, B, _i(out))
        return out


def read_idx_images(path, max_rows=1 << 30):
    import struct
    with open(path, "rb") as f:
        _, n, h, w = struct.unpack(">iiii", f.read(16))
    n = min(n, max_rows)
    out = np.empty((n, h * w), dtype=np.float64)
    rl = C.c_int32(0)
    got = lib().orc_read_idx_images(path.encode(), _d(out), n, C.byref(rl))
    if got < 0:
        raise IOError("orc_read_idx_images(%s) -> %d" % (path, got))
    return out[:got]


def read_idx_labels(path, max_rows=1 << 30):
    import struct
    with open(path, "rb") as f:
        _, n = struct.unpack(">ii", f.read(8))
    n = min(n, max_rows)
    out = np.empty(n, dtype=np.int32)
    got = lib().orc_read_idx_labels(path.encode(), _i(out), n)
    if got < 0:
        raise IOError("orc_read_idx_labels(%s) -> %d" % (path, got))
    return out[:got]


def format_log_row(dims, iterations, step, batch, momentum, noise, train_acc, test_acc):
    buf = C.create_string_buffer(512)
    d = np.asarray(dims, dtype=np.int32)
    n = lib().orc_format_log_row(buf, 512, _i(d), len(dims), iterations, step, batch, momentum,
                                 int(noise), train_acc, test_acc)
    return buf.raw[:n].decode()
