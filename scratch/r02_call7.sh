#!/bin/bash
# 1 GPU: epilogue-warp variants of the GEMM (4 / 8 / 16 warps), stall counters of each, then the round's ncu evidence
mkdir -p gpurun_out/r02g
O=gpurun_out/r02g
L=graph-neural-net_b200/lib
for v in "" _e16 _e4; do
  B2NN_LIB=$PWD/$L/libb2nn$v.so python bench.py --no-cpu --no-also --no-sustained > $O/bench$v.json 2> $O/bench$v.err; echo "bench$v rc=$?"
  B2NN_LIB=$PWD/$L/libb2nn$v.so B2NN_GEMM_STALLS=1 python scratch/stalls.py 2>&1 | tail -8 > $O/stalls$v.log
done
python - <<'PY'
import json
for v in ("","_e16","_e4"):
    try:
        d=json.load(open("gpurun_out/r02g/bench%s.json"%v)); r=d["roofline"]
        print("epi warps %s: ms/step %.4f gemm avg %.4f serialised %.4f e2e %.3f"%(v or "_e8", d["ms_per_step"], r["avg_launch_ms"], r["serialised_step_ms"], d["e2e"]["ms_per_step"]))
    except Exception as e: print(v, "failed", e)
PY
for v in "" _e16 _e4; do echo "== stalls$v"; cat $O/stalls$v.log | cut -c1-330; done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-also --no-sustained"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4.csv $CMD > $O/ncu_1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 24 -c 8 -o $O/c4_gemm $CMD > $O/ncu_2.log 2>&1
echo ncu rc=$?
CMD2="python bench.py --workload c2 --steps 300 --warmup 3 --no-cpu"
$CMD2 > $O/plain_c2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fused_train -c 2 -o $O/c2_k1 $CMD2 > $O/ncu_3.log 2>&1
echo ncu c2 rc=$?
ls -la $O/*.ncu-rep
