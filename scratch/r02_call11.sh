#!/bin/bash
mkdir -p gpurun_out/r02k
O=gpurun_out/r02k
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py tests/test_gpu_zz_checkpoint.py -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -3 $O/pytest_gpu.log
python bench.py --no-cpu --no-also > $O/bench_fused.json 2> $O/bench_fused.err; echo rc=$?
python - <<'PY'
import json
for v in ("fused",):
    d=json.load(open("gpurun_out/r02k/bench_%s.json"%v)); r=d["roofline"]
    print("%s: ms/step %.4f (step TF %.0f) gemm avg %.4f achieved %.0f serialised %.4f e2e %.3f"%(v, d["ms_per_step"], d["step_tflops_per_gpu"], r["avg_launch_ms"], r["achieved"], r["serialised_step_ms"], d["e2e"]["ms_per_step"]), r["non_gemm_us"], d.get("sustained"), "launches", d["gpu_launches"])
PY
B2NN_GEMM_STALLS=1 python scratch/stalls.py 2>&1 | tail -8 | cut -c1-330
