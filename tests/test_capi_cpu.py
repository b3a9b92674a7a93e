"""CPU-side tests of the boundary: libb2nn.so loads without a GPU, exports every symbol
include/b2nn.h declares, reports NO_DEVICE instead of falling back to a CPU path, and the host
logic above the C-ABI (function singletons, sampler, MNIST harness pieces) behaves like the
reference's.  No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "b2nn.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2nn_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(g):
    names = _declared_symbols()
    assert len(names) >= 25
    lib = C.CDLL(g.capi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libb2nn.so does not export %s" % n
    # and the Python binding covers exactly the header
    assert sorted(g.capi.SIGNATURES) == names


def test_version_and_error_text(g):
    assert g.capi.lib().b2nn_version() == 100
    assert isinstance(g.capi.lib().b2nn_last_error(), bytes)


def test_no_cpu_fallback_without_device(g):
    if g.capi.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(g.capi.B2nnError) as e:
        g.SoftmaxCrossEntropyNeuralNet([4, 3, 2], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    assert e.value.code == g.capi.ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_constructor_argument_errors(g):
    # argument validation happens before any device is touched
    with pytest.raises(ValueError):
        g.SoftmaxCrossEntropyNeuralNet([4], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)          # S:105
    with pytest.raises(ValueError):
        g.SoftmaxCrossEntropyNeuralNet([4, 0, 2], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)    # S:141
    with pytest.raises(ValueError):
        g.SoftmaxCrossEntropyNeuralNet([4, 2], None, g.LEAKY_RELU_PRIME)               # S:106
    with pytest.raises(ValueError):                                                   # arbitrary lambda: no kernel
        g.SoftmaxCrossEntropyNeuralNet([4, 2], lambda a: a, lambda a: 1.0)
    with pytest.raises(ValueError):                                                   # mismatched pair
        g.SoftmaxCrossEntropyNeuralNet([4, 2], g.LEAKY_RELU, g.SIGMOID_PRIME)
    with pytest.raises(ValueError):
        g.GeneralNeuralNet([4, 2], g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME,
                           lambda c, e: 0.0, lambda c, e: 0.0)
    with pytest.raises(ValueError):                                                   # unknown precision enum
        g.SoftmaxCrossEntropyNeuralNet([4, 2], g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=7)
    with pytest.raises(g.capi.B2nnError) as e:                                        # TF32 is a real mode: needs a GPU
        g.SoftmaxCrossEntropyNeuralNet([4, 2], g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_TF32)
    assert e.value.code == g.capi.ERR_NO_DEVICE


def test_function_singletons_match_oracle_scalars(g, orc):
    zs = [-3.0, -0.5, 0.0, 1e-9, 0.7, 4.0]
    pairs = [("leaky_relu", 0, g.LEAKY_RELU, g.LEAKY_RELU_PRIME, 0.01), ("sigmoid", 1, g.SIGMOID, g.SIGMOID_PRIME, 0.0),
             ("relu", 2, g.RELU, g.RELU_PRIME, 0.0), ("tanh", 3, g.TANH, g.TANH_PRIME, 0.0),
             ("identity", 4, g.IDENTITY, g.IDENTITY_PRIME, 0.0)]
    for name, kid, f, fp, p in pairs:
        assert f.kernel_id == kid == orc.ACT[name] and fp.kernel_id == kid
        for z in zs:
            assert abs(f.func(z) - orc.lib().orc_act(kid, p, z)) <= 1e-15
            assert abs(fp.func(z) - orc.lib().orc_act_prime(kid, p, z)) <= 1e-15
    assert g.HALF_SQUARED_ERROR.loss(0.8, 1.0) == orc.lib().orc_loss(0, 0.8, 1.0)
    assert g.HALF_SQUARED_ERROR_PRIME.func(0.8, 1.0) == orc.lib().orc_loss_prime(0, 0.8, 1.0)
    # the in-tree lambdas (MNISTTrainer.java:234-235)
    assert g.LEAKY_RELU.func(-2.0) == 0.01 * -2.0 and g.LEAKY_RELU_PRIME.func(0.0) == 0.01


@pytest.mark.parametrize("n,B,iters", [(60000, 16, 7600), (100, 7, 40), (17, 16, 9)])
def test_native_sampler_matches_oracle(g, orc, n, B, iters):
    got = g.Sampler(n, seed=1).draw(iters, B)
    s = orc.Sampler(n, seed=1)
    ref = np.stack([s.sample(B) for _ in range(iters)])
    assert (got == ref).all()


def test_sampler_argument_errors(g):
    with pytest.raises(ValueError):
        g.Sampler(0)
    with pytest.raises(ValueError):
        g.Sampler(10).draw(1, 10)   # batchSize < rows is strict (T:63, Q17)


def test_log_row_and_idx_readers_match_oracle(g, orc, golden_dir, tmp_path):
    from gnn_b200 import mnist
    row = mnist.MNISTTrainer.formatLogRow([784, 300, 100, 10], 100000, 0.0125, 16, 0.9, False, 0.99973, 0.978)
    assert row == orc.format_log_row([784, 300, 100, 10], 100000, 0.0125, 16, 0.9, False, 0.99973, 0.978)
    assert row == open(os.path.join(golden_dir, "trainLog_reference.csv")).read().splitlines()[3] + "\n"
    lp = os.path.join(golden_dir, "mnist", "t10k-labels-idx1-ubyte")
    assert (mnist.read_idx_labels(lp) == orc.read_idx_labels(lp)).all()
    ip = str(tmp_path / "img")
    mnist.write_surrogate_images(mnist.read_idx_labels(lp)[:50], ip, seed=2)
    a, b = mnist.read_idx_images(ip), orc.read_idx_images(ip)
    assert a.shape == (50, 784) and (a == b).all() and a.min() >= 0.0 and a.max() <= 1.0
    ip2 = str(tmp_path / "img2")
    mnist.write_surrogate_images(mnist.read_idx_labels(lp)[:50], ip2, seed=2)
    assert open(ip, "rb").read() == open(ip2, "rb").read()   # deterministic
    oh = mnist.one_hot(np.array([3, 0, 9]))
    assert oh.shape == (3, 10) and oh[0, 3] == 1 and oh.sum() == 3


def test_pixel_byte_quotient_float_equals_double():
    """the wide path reads B2NN_DT_U8 rows as the IEEE float quotient byte / 255.0f; for all 256 bytes this is the
    double quotient of the reference's loader (MNISTTrainer.java:97) rounded to float, so byte rows and converted FP32
    rows feed the kernels identical values"""
    b = np.arange(256)
    assert np.array_equal((b.astype(np.float32) / np.float32(255.0)).astype(np.float32),
                          (b.astype(np.float64) / 255.0).astype(np.float32))


@pytest.mark.parametrize("M,N,pairs", [(4096, 4096, 74), (4096, 1000, 74), (32768, 8192, 74), (8192, 8192, 74),
                                        (300, 264, 74), (37, 8, 74), (2560, 2048, 74), (4096, 4096, 8), (2304, 512, 5),
                                        (256 * 9, 256 * 3, 74), (256 * 17, 256 * 5, 37), (1, 1, 1)])
def test_tile_order_covers_every_output_once(g, M, N, pairs):
    """The contraction kernel's work list (TileSched in gemm_tcgen05.cuh, decoded on the host through
    b2nn_debug_tile_list): bands of 8 tile rows walked column by column, the last partial wave split into 256 x 128
    halves.  Every 256 x 256 output tile must appear exactly once -- whole, or as its two halves -- whatever the
    shape (ragged edges, fewer than 8 tile rows, a short last band), and the items a wave of `pairs` pairs holds at a time
    must share operand slabs the way the order promises."""
    lib = g.capi.lib()
    total = C.c_int32()
    g.capi.check(lib.b2nn_debug_tile_list(M, N, pairs, None, None, None, 0, C.byref(total)))
    n = total.value
    m0, n0, bn = (np.empty(n, np.int32) for _ in range(3))
    ip = C.POINTER(C.c_int32)
    g.capi.check(lib.b2nn_debug_tile_list(M, N, pairs, m0.ctypes.data_as(ip), n0.ctypes.data_as(ip),
                                          bn.ctypes.data_as(ip), n, C.byref(total)))
    m_tiles, n_tiles = -(-M // 256), -(-N // 256)
    cover = np.zeros((m_tiles, 2 * n_tiles), np.int32)          # in units of 128 columns
    for a, b, w in zip(m0, n0, bn):
        assert a % 256 == 0 and b % 128 == 0 and w in (128, 256) and 0 <= a < m_tiles * 256 and 0 <= b < n_tiles * 256
        assert w == 128 or b % 256 == 0
        cover[a // 256, b // 128: b // 128 + w // 128] += 1
    assert (cover == 1).all()
    T = m_tiles * n_tiles
    rem = T % pairs
    halves = int((bn == 128).sum())
    assert halves == (2 * rem if 0 < rem and 2 * rem <= pairs else 0) and n == T + halves // 2
    assert (bn[:n - halves] == 256).all()                        # the halves come last, in pairs of one tile
    for i in range(n - halves, n, 2):
        assert m0[i] == m0[i + 1] and n0[i] + 128 == n0[i + 1]
    if m_tiles >= 8 and n_tiles >= 10 and pairs >= 64:           # the first wave is an 8-row patch, not 2-3 whole rows
        first = slice(0, pairs)
        assert len(set(m0[first].tolist())) == 8 and len(set(n0[first].tolist())) <= -(-pairs // 8) + 1


def test_reference_arm_contract_line():
    """`bench.py --impl reference` (the driver's reference arm): rank 0 prints ONE JSON line with the contract's keys, timed on
    the host by the oracle's faithful loops; every other rank of a torchrun launch exits 0 without output or work."""
    import json
    import subprocess
    import sys
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1", "--steps", "3",
           "--warmup", "3", "--gpus", "2"]
    env = dict(os.environ, RANK="0", WORLD_SIZE="2", LOCAL_RANK="0")
    r0 = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r0.returncode == 0, r0.stderr[-2000:]
    lines = [l for l in r0.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "train_samples_per_sec" and d["unit"] == "samples/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 2 and d["steps"] == 3 and d["warmup"] == 3
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("c1")
    r1 = subprocess.run(cmd, env=dict(env, RANK="1", LOCAL_RANK="1"), capture_output=True, text=True, timeout=300)
    assert r1.returncode == 0 and r1.stdout.strip() == ""
