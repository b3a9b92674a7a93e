"""Host-side checks of the checkpoint file (graph-neural-net_b200/csrc/checkpoint.cu) that need no GPU: the header
reader against a file assembled here from the documented byte layout, and the error paths (bad magic, version,
truncation, dimension / weight-count mismatch, null arguments).  Save / load round trips are GPU tests
(tests/test_gpu_z_checkpoint.py)."""
import ctypes as C
import struct

import numpy as np
import pytest


def fnv1a64(data):
    h = 0xcbf29ce484222325
    for b in data:
        h ^= b
        h = (h * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h


def build_file(dims, w, v, time=7, kind=0, inner_act=0, last_act=0, loss=0, precision=0, flags=0, inner_param=0.01,
               last_param=0.0, version=1, magic=b"B2NNCKPT", n_weights=None):
    """the layout documented at the top of checkpoint.cu, little-endian, no padding"""
    nw = len(w) if n_weights is None else n_weights
    body = magic + struct.pack("<II", version, len(dims)) + struct.pack("<%di" % len(dims), *dims)
    body += struct.pack("<6i", kind, inner_act, last_act, loss, precision, flags)
    body += struct.pack("<2d", inner_param, last_param) + struct.pack("<2q", time, nw)
    body += np.asarray(w, dtype="<f8").tobytes() + np.asarray(v, dtype="<f8").tobytes()
    return body + struct.pack("<Q", fnv1a64(body))


def test_info_reads_the_documented_layout(g, tmp_path):
    dims = [5, 4, 3]
    n = 5 * 4 + 4 * 3
    p = tmp_path / "a.ckpt"
    p.write_bytes(build_file(dims, np.arange(n), np.zeros(n), time=123, kind=1, precision=2))
    info = g.capi.checkpoint_info(str(p))
    assert info == {"layerDims": dims, "kind": 1, "precision": 2, "time": 123, "numWeights": n}


def test_info_error_paths(g, tmp_path):
    dims, n = [5, 4, 3], 32
    good = build_file(dims, np.arange(n), np.zeros(n))
    cases = {
        "magic": (build_file(dims, np.arange(n), np.zeros(n), magic=b"NOTACKPT"), ValueError),
        "version": (build_file(dims, np.arange(n), np.zeros(n), version=9), NotImplementedError),
        "count": (build_file(dims, np.arange(n), np.zeros(n), n_weights=n + 1), ValueError),
        "zero_dim": (build_file([5, 0, 3], [], []), ValueError),
        "truncated": (good[:30], ValueError),
        "empty": (b"", ValueError),
    }
    for name, (blob, exc) in cases.items():
        p = tmp_path / (name + ".ckpt")
        p.write_bytes(blob)
        with pytest.raises(exc):
            g.capi.checkpoint_info(str(p))
    with pytest.raises(ValueError):
        g.capi.checkpoint_info(str(tmp_path / "does_not_exist.ckpt"))
    # a dims buffer that is too small is refused, not overrun
    p = tmp_path / "good.ckpt"
    p.write_bytes(good)
    small = (C.c_int32 * 2)()
    rc = g.capi.lib().b2nn_checkpoint_info(str(p).encode(), None, small, 2, None, None, None, None)
    assert rc == g.capi.ERR_INVALID_ARG


def test_null_arguments(g):
    L = g.capi.lib()
    assert L.b2nn_checkpoint_save(None, b"/tmp/x") == g.capi.ERR_INVALID_ARG
    assert L.b2nn_checkpoint_load(None, b"/tmp/x") == g.capi.ERR_INVALID_ARG
    assert L.b2nn_checkpoint_info(None, None, None, 0, None, None, None, None) == g.capi.ERR_INVALID_ARG
    assert L.b2nn_time(None, None) == g.capi.ERR_INVALID_ARG
