#!/bin/bash
# 2 GPUs: data-parallel parity tests (bf16 / tf32 x p2p / nccl / nvls at world 2), bench with dp_check, one timeline
mkdir -p gpurun_out/r02n2
O=gpurun_out/r02n2
timeout 900 python -m pytest tests/test_dp.py -q -m gpu -k "world_matches" > $O/pytest_dp.log 2>&1; echo pytest rc=$?; tail -15 $O/pytest_dp.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --no-cpu > $O/bench_n2.json 2> $O/bench_n2.err; echo bench rc=$?
tail -3 $O/bench_n2.err
B2NN_DP=nvls python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --no-cpu --no-also --no-sustained --precision tf32 > $O/bench_n2_tf32_nvls.json 2> $O/bench_n2_tf32_nvls.err; echo bench tf32 rc=$?
B2NN_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 2 --steps 8 --no-cpu --no-also --no-sustained > $O/bench_n2_trace.json 2> $O/bench_n2_trace.err
python - <<'PY'
import json
for f in ("bench_n2","bench_n2_tf32_nvls"):
    try:
        d=json.load(open("gpurun_out/r02n2/%s.json"%f))
        print(f, "ms/step %.4f value %.3g"%(d["ms_per_step"], d["value"]), d["config"]["exchange"], d.get("dp_check"), "e2e %.3f"%d["e2e"]["ms_per_step"], d.get("sustained"))
        for k,v in d.get("also",{}).items():
            print("  also", k, {q: v.get(q) for q in ("ms_per_step","value","batch_per_gpu","exchange","error")})
    except Exception as e:
        print(f, "failed", e)
PY
grep "trace r0" $O/bench_n2_trace.err | head -60
