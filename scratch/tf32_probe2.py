"""TF32 bring-up probe 2: sweep the source-format fields of the instruction descriptor on all-ones operands.
B2NN_GEMM_DBG is read once per process now: run one process per format (B2NN_GEMM_DBG = 1 | fmt << 4)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
g = ge.load_package()
M, K, N = 128, 64, 256
net = g.GeneralNeuralNet([K, N], g.IDENTITY, g.IDENTITY_PRIME, g.IDENTITY, g.IDENTITY_PRIME, g.HALF_SQUARED_ERROR,
                         g.HALF_SQUARED_ERROR_PRIME, precision=g.capi.PREC_TF32)
net.setWeights(np.ones(K * N))
x = np.ones((M, K))
for fmt in [None, 0, 1, 2, 3, 4, 5, 6, 7]:
    if fmt is None:
        os.environ.pop("B2NN_GEMM_DBG", None)
    else:
        os.environ["B2NN_GEMM_DBG"] = str(1 | (fmt << 4))
    try:
        y = net.propagate(x)
        print("fmt", fmt, "min %.5g max %.5g mean %.5g nan %d  y[0,:6]=%s y[127,250:]=%s" %
              (np.nanmin(y), np.nanmax(y), np.nanmean(y), np.isnan(y).sum(), y[0, :6], y[127, 250:]), flush=True)
    except Exception as e:
        print("fmt", fmt, "ERROR", repr(e)[:200], flush=True)
        break
