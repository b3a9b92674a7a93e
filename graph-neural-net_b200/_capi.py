"""ctypes binding of libb2nn.so (include/b2nn.h).  This is the same surface the Java shim binds
through JNI / Panama FFM; nothing here computes anything -- if the CUDA library
is missing the import of the product fails loudly (there is no CPU fallback)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2NN_LIB selects a measurement variant of the library (build.py --variant); the product is lib/libb2nn.so
LIB_PATH = os.environ.get("B2NN_LIB") or os.path.join(_HERE, "lib", "libb2nn.so")

OK = 0
ERR_INVALID_ARG, ERR_DIM_MISMATCH, ERR_UNSUPPORTED, ERR_CUDA, ERR_NCCL, ERR_OOM, ERR_NO_DEVICE = -1, -2, -3, -4, -5, -6, -7
NET_SOFTMAX_CE, NET_GENERAL = 0, 1
PREC_F64, PREC_BF16, PREC_TF32 = 0, 1, 2
FLAG_SCALED_INIT, FLAG_STABLE_SOFTMAX, FLAG_BF16_WIRE = 1, 2, 4
DT_F64, DT_F32, DT_BF16, DT_U8 = 0, 1, 2, 3
DATA_DEVICE, DATA_HOST_PINNED = 0, 1

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)
_vp = C.c_void_p

# every symbol include/b2nn.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "b2nn_create": (C.c_int, [_ip, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_int32, C.c_double, C.c_int32,
                              C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.POINTER(_vp)]),
    "b2nn_destroy": (C.c_int, [_vp]),
    "b2nn_dims": (C.c_int, [_vp, _ip, _ip]),
    "b2nn_num_weights": (C.c_int, [_vp, _lp]),
    "b2nn_propagate": (C.c_int, [_vp, _dp, C.c_int32, _dp]),
    "b2nn_loss": (C.c_int, [_vp, _dp, _dp, C.c_int32, _dp]),
    "b2nn_weight_gradient": (C.c_int, [_vp, _dp, _dp, _dp]),
    "b2nn_gradient_step": (C.c_int, [_vp, _dp, _dp, C.c_int32, C.c_double, C.c_double, C.c_int32, _dp]),
    "b2nn_gradient_step_labeled": (C.c_int, [_vp, _vp, C.c_int32, _ip, C.c_int32, C.c_double, C.c_double, _dp]),
    "b2nn_get_weights": (C.c_int, [_vp, _dp]),
    "b2nn_set_weights": (C.c_int, [_vp, _dp]),
    "b2nn_get_velocity": (C.c_int, [_vp, _dp]),
    "b2nn_set_velocity": (C.c_int, [_vp, _dp]),
    "b2nn_time": (C.c_int, [_vp, _lp]),
    "b2nn_checkpoint_save": (C.c_int, [_vp, C.c_char_p]),
    "b2nn_checkpoint_load": (C.c_int, [_vp, C.c_char_p]),
    "b2nn_checkpoint_info": (C.c_int, [C.c_char_p, _ip, _ip, C.c_int32, _ip, _ip, _lp, _lp]),
    "b2nn_dataset_create": (C.c_int, [_vp, _vp, C.c_int32, _dp, _ip, C.c_int64, C.c_int32, C.POINTER(_vp)]),
    "b2nn_dataset_destroy": (C.c_int, [_vp]),
    "b2nn_train_indexed": (C.c_int, [_vp, _vp, _ip, C.c_int32, C.c_int32, C.c_double, C.c_double, _dp]),
    "b2nn_validate": (C.c_int, [_vp, _vp, C.c_int32, _dp]),
    "b2nn_evaluate": (C.c_int, [_vp, _vp, C.c_int64, _lp, _dp, _ip]),
    "b2nn_sampler_create": (C.c_int, [C.c_int32, C.c_int64, C.POINTER(_vp)]),
    "b2nn_sampler_draw": (C.c_int, [_vp, C.c_int32, C.c_int32, _ip]),
    "b2nn_sampler_destroy": (C.c_int, [_vp]),
    "b2nn_comm_unique_id": (C.c_int, [_vp]),
    "b2nn_comm_init": (C.c_int, [_vp, _vp, C.c_int32, C.c_int32]),
    "b2nn_comm_ipc_export": (C.c_int, [_vp, _vp]),
    "b2nn_comm_ipc_import": (C.c_int, [_vp, _vp, C.c_int32]),
    "b2nn_comm_nvls_export": (C.c_int, [_vp, _vp]),
    "b2nn_comm_nvls_import": (C.c_int, [_vp, _vp, C.c_int32]),
    "b2nn_comm_nvls_team": (C.c_int, [_vp, _vp]),
    "b2nn_comm_nvls_bind": (C.c_int, [_vp]),
    "b2nn_comm_mode": (C.c_int, [_vp, _vp]),
    "b2nn_sync": (C.c_int, [_vp]),
    "b2nn_propagate_device_noalloc": (C.c_int, [_vp, _vp, C.c_int32, _vp]),   # DEVICE pointers
    "b2nn_noalloc_max_doubles": (C.c_int, [C.POINTER(C.c_int64)]),
    "b2nn_set_stream": (C.c_int, [_vp, _vp]),
    "b2nn_kernel_launches": (C.c_int, [_vp, _lp]),
    "b2nn_profile_enable": (C.c_int, [_vp, C.c_int32]),
    "b2nn_profile_read": (C.c_int, [_vp, _dp, _lp, _dp]),
    "b2nn_profile_read_classes": (C.c_int, [_vp, _dp, _lp]),
    "b2nn_profile_read_stamps": (C.c_int, [_vp, _dp, _dp, C.POINTER(C.c_int64), _dp]),
    "b2nn_profile_stamp_clock_mhz": (C.c_int, [_vp, _dp]),
    "b2nn_k1_profile": (C.c_int, [_vp, C.c_int32, _dp, _ip]),
    "b2nn_debug_tile_list": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, _ip, _ip, _ip, C.c_int32, _ip]),
    "b2nn_device_count": (C.c_int, [_ip]),
    "b2nn_version": (C.c_int, []),
    "b2nn_last_error": (C.c_char_p, []),
}

_lib = None


def lib():
    """Load libb2nn.so (raises OSError if it has not been built: run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError("libb2nn.so not built at %s: run `python graph-neural-net_b200/build.py`; "
                          "this backend has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


class B2nnError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b2nn error %d: %s" % (code, msg))
        self.code = code


def check(code):
    """Map B2NN_ERR_* like the Java shim does: INVALID_ARG/DIM_MISMATCH -> ValueError
    (IllegalArgumentException), UNSUPPORTED -> NotImplementedError (UnsupportedOperationException),
    the rest -> B2nnError (IllegalStateException)."""
    if code == OK:
        return
    msg = (lib().b2nn_last_error() or b"").decode(errors="replace")
    if code in (ERR_INVALID_ARG, ERR_DIM_MISMATCH):
        raise ValueError("b2nn: " + msg)
    if code == ERR_UNSUPPORTED:
        raise NotImplementedError("b2nn: " + msg)
    raise B2nnError(code, msg)


def checkpoint_info(path):
    """header of a checkpoint file (b2nn_checkpoint_info; host only): dict with layerDims, kind, precision, time,
    numWeights"""
    import numpy as np
    nd, kind, prec = C.c_int32(0), C.c_int32(0), C.c_int32(0)
    t, nw = C.c_int64(0), C.c_int64(0)
    p = os.fsencode(path)
    check(lib().b2nn_checkpoint_info(p, C.byref(nd), None, 0, C.byref(kind), C.byref(prec), C.byref(t), C.byref(nw)))
    dims = np.zeros(nd.value, dtype=np.int32)
    check(lib().b2nn_checkpoint_info(p, None, dims.ctypes.data_as(_ip), nd.value, None, None, None, None))
    return {"layerDims": [int(d) for d in dims], "kind": kind.value, "precision": prec.value, "time": t.value,
            "numWeights": nw.value}


def device_count():
    n = C.c_int32(0)
    check(lib().b2nn_device_count(C.byref(n)))
    return n.value
