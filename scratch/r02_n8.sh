#!/bin/bash
# 8 GPUs: the driver's SCALE command (default bench: C4 weak + dp_check + also.c5 strong), TF32 with the in-switch exchange,
# one timeline, then the in-switch parity tests at world 8 / 4
mkdir -p gpurun_out/r02n8
O=gpurun_out/r02n8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 8 --no-cpu > $O/bench_n8.json 2> $O/bench_n8.err; echo bench rc=$?
tail -2 $O/bench_n8.err
$TR --master-port 29522 bench.py --gpus 8 --no-cpu --no-also --no-sustained --precision tf32 --steps 20 > $O/bench_n8_tf32.json 2> $O/bench_n8_tf32.err; echo bench tf32 rc=$?
B2NN_TRACE=1 $TR --master-port 29523 bench.py --gpus 8 --steps 8 --no-cpu --no-also --no-sustained > $O/bench_n8_trace.json 2> $O/bench_n8_trace.err
python - <<'PY'
import json
for f in ("bench_n8","bench_n8_tf32"):
    try:
        d=json.load(open("gpurun_out/r02n8/%s.json"%f))
        print(f, "ms/step %.4f value %.4g"%(d["ms_per_step"], d["value"]), d["config"]["exchange"], d.get("dp_check"), "e2e %.3f"%d["e2e"]["ms_per_step"], d.get("sustained"))
        for k,v in d.get("also",{}).items():
            print("  also", k, {q: v.get(q) for q in ("ms_per_step","value","batch_per_gpu","exchange","step_tflops_per_gpu","error")})
    except Exception as e:
        print(f, "failed", e)
PY
grep "trace r0" $O/bench_n8_trace.err | head -36
timeout 400 python -m pytest tests/test_dp.py -q -m gpu -k "(bf16-nvls-8 or tf32-nvls-8 or bf16-nvls-4 or bf16-p2p-8)" > $O/pytest_dp.log 2>&1; echo pytest rc=$?; tail -6 $O/pytest_dp.log
