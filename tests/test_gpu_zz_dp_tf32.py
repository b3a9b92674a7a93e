"""2-GPU test of data-parallel TF32 training (skipped on a 1-GPU box).  Sorted last and bounded in time on purpose:
it was written after round 1's GPU budget was spent and has not run on two B200s yet."""
import socket
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_world2_tf32_matches_single_gpu(g, tmp_path, monkeypatch):
    """TF32 nets shard the batch like BF16 nets (tests/test_dp.py); dp.init_comm gives them the NCCL all-reduce +
    local update exchange whatever B2NN_DP says (they have no BF16 operand copy for the fused kernel to replicate).
    Replicas bit-identical; weights equal to the single-GPU run within 2e-3 of the update's Frobenius norm over 6
    steps (only the FP32 summation order of the batch sum differs)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from tests import dp_workers
    monkeypatch.setenv("B2NN_DP", "p2p")
    ctx = mp.spawn(dp_workers.nccl_worker, args=(2, _free_port(), str(tmp_path), "tf32"), nprocs=2, join=False)
    deadline = time.time() + 180
    while not ctx.join(timeout=5):
        if time.time() > deadline:
            for p in ctx.processes:
                p.kill()
            pytest.fail("2-rank TF32 run did not finish within 180 s")
    r = np.load(tmp_path / "nccl.npz")
    assert int(r["replicas_equal"]) == 1
    du_dp, du_1 = r["w_dp"] - r["w0"], r["w_single"] - r["w0"]
    assert np.linalg.norm(du_dp - du_1) / np.linalg.norm(du_1) < 2e-3
    assert str(r["mode"]) == "nccl"
