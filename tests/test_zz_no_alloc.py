"""b2nn_propagate_device_noalloc (include/b2nn.h; no counterpart in the reference):
the FP64 forward pass over DEVICE buffers with its work rows in a fixed __device__ array, for callers that
may not allocate.  Checked through the C-ABI against the existing entry point b2nn_propagate on the same net and rows,
two calls in a row (the second call finds the first call's contents in the shared array), at the tolerance of the FP64
parity tests (1e-13; the kernels and the summation order are the same, so the outputs are in fact equal bit for bit).

Collected last on purpose (pytest -x must not lose the other tests to a failure here); passes alone in
a fresh process and after the other tests.  The TEST allocates the caller's buffers with torch; the entry point does not
allocate (free device memory is compared across the second call).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-13


def _nets(g):
    yield "softmax", g.SoftmaxCrossEntropyNeuralNet([784, 300, 100, 10], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    yield "general", g.GeneralNeuralNet([50, 33, 7], g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME,
                                        g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME)


@pytest.mark.parametrize("rows", [(16, 16), (200, 7)])     # <= 32 rows and > 32 rows pick different tile shapes
def test_noalloc_propagate_matches_propagate_twice(g, rows):
    import torch
    dev = torch.device("cuda:0")
    for name, net in _nets(g):
        d0, dL = net.layerDims[0], net.layerDims[-1]
        rng = np.random.RandomState(5)
        for call, n in enumerate(rows):                     # two calls in a row, the second on other rows
            x = rng.rand(n, d0)
            want = net.propagate(x)                         # existing entry point: host rows, cudaMalloc'ed work buffers
            x_dev = torch.from_numpy(x).to(dev)
            y_dev = torch.full((n, dL), float("nan"), dtype=torch.float64, device=dev)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info(0)[0]
            net.propagateDeviceNoAlloc(x_dev.data_ptr(), n, y_dev.data_ptr())
            net.sync()
            free1 = torch.cuda.mem_get_info(0)[0]
            got = y_dev.cpu().numpy()
            assert np.isfinite(got).all(), (name, call)
            assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max()), (name, call, np.abs(got - want).max())
            if call == 1:                                   # (the first call may load the kernels' module lazily)
                assert free1 == free0, "device memory changed across an allocation-free call: %d -> %d" % (free0, free1)
        net.close()


def test_noalloc_rejects_what_it_cannot_hold(g):
    import ctypes
    import torch
    cap = ctypes.c_int64(0)
    g.capi.check(g.capi.lib().b2nn_noalloc_max_doubles(ctypes.byref(cap)))
    assert cap.value == 524288                              # B2NN_NOALLOC_MAX_DOUBLES: 4 MB
    net = g.SoftmaxCrossEntropyNeuralNet([784, 300, 100, 10], g.LEAKY_RELU, g.LEAKY_RELU_PRIME)
    per_row = 784 + 2 * (300 + 100) + 10
    n = cap.value // per_row + 64
    buf = torch.zeros(n * 784, dtype=torch.float64, device="cuda:0")
    out = torch.zeros(n * 10, dtype=torch.float64, device="cuda:0")
    with pytest.raises(ValueError, match="fixed scratch array"):
        net.propagateDeviceNoAlloc(buf.data_ptr(), n, out.data_ptr())
    with pytest.raises(ValueError):
        net.propagateDeviceNoAlloc(0, 4, out.data_ptr())
    net.propagateDeviceNoAlloc(buf.data_ptr(), 8, out.data_ptr())   # still serves a size that fits
    net.sync()
    net.close()
    wide = g.SoftmaxCrossEntropyNeuralNet([256, 128, 16], g.LEAKY_RELU, g.LEAKY_RELU_PRIME, precision=g.capi.PREC_BF16,
                                          flags=g.capi.FLAG_SCALED_INIT | g.capi.FLAG_STABLE_SOFTMAX)
    with pytest.raises(NotImplementedError):
        wide.propagateDeviceNoAlloc(buf.data_ptr(), 4, out.data_ptr())
    wide.close()
