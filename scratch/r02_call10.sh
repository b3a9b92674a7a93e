#!/bin/bash
# fused-update epilogue (default on 1 GPU) against the separate update kernel, 5 stages: parity, numbers, stall counters
mkdir -p gpurun_out/r02j
O=gpurun_out/r02j
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -5 $O/pytest_gpu.log
python bench.py --no-cpu --no-also > $O/bench_fused.json 2> $O/bench_fused.err; echo rc=$?
B2NN_FUSED_UPDATE=0 python bench.py --no-cpu --no-also > $O/bench_unfused.json 2> $O/bench_unfused.err; echo rc=$?
python bench.py --no-cpu --no-also --no-sustained --precision tf32 > $O/bench_fused_tf32.json 2> $O/bench_fused_tf32.err; echo rc=$?
python - <<'PY'
import json
for v in ("fused","unfused","fused_tf32"):
    try:
        d=json.load(open("gpurun_out/r02j/bench_%s.json"%v)); r=d["roofline"]
        print("%s: ms/step %.4f (step TF %.0f) gemm avg %.4f achieved %.0f serialised %.4f e2e %.3f"%(v, d["ms_per_step"], d["step_tflops_per_gpu"], r["avg_launch_ms"], r["achieved"], r["serialised_step_ms"], d["e2e"]["ms_per_step"]), r["non_gemm_us"], d.get("sustained"), "launches", d["gpu_launches"])
    except Exception as e: print(v, "failed", e)
PY
B2NN_GEMM_STALLS=1 python scratch/stalls.py 2>&1 | tail -8 | cut -c1-330
python scratch/tf32_batch_probe.py 2>&1 | tail -12
B2NN_FUSED_UPDATE=0 python scratch/tf32_batch_probe.py 2>&1 | tail -12
