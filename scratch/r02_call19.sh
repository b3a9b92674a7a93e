#!/bin/bash
# round 2, 1 GPU: band height of the tile order (6 / 8 / 12 / 16) and evict_last on the band's row slabs, on C5 at its
# 1-GPU shape (32768 rows) and at the per-GPU shape of the 8-GPU step (4096 rows); DRAM bytes per launch for two of them
#   gpurun --timeout 1200 -- 'bash scratch/r02_call19.sh'
mkdir -p gpurun_out/r02v
O=gpurun_out/r02v
L=graph-neural-net_b200/lib
show() { python -c "
import json,sys; d=json.load(open('$1')); r=d['roofline']; k=r['in_kernel_timer']
print('$2: ms/step %.4f'%d['ms_per_step'], 'busy %.4f'%k['contraction_busy_ms_per_step'], 'eff MHz %.0f'%k['sm_mhz_effective'], 'frac_burst %.3f'%r['frac_burst'])"; }
for v in base def band6 band12 band16 hintA band16hintA def; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5_$v.json 2> $O/bench_c5_$v.err; show $O/bench_c5_$v.json "c5 32768 rows $v"
done
export B2NN_BENCH_C5_ROWS=4096
for v in base def band6 band12 band16 hintA band16hintA def; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --workload c5 --steps 20 --warmup 5 --no-cpu --no-also --no-sustained > $O/bench_c5r4096_$v.json 2> $O/bench_c5r4096_$v.err; show $O/bench_c5r4096_$v.json "c5 4096 rows $v"
done
unset B2NN_BENCH_C5_ROWS
for v in hintA band16hintA; do
  export B2NN_LIB=$L/libb2nn_$v.so
  timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_bf16 -s 20 -c 2 --csv --log-file $O/ncu_c5_$v.csv python bench.py --workload c5 --steps 2 --warmup 1 --no-cpu --no-also --no-sustained > $O/ncu_c5_$v.log 2>&1; echo "ncu $v rc=$?"
  grep -E "dram__bytes_read|gpu__time" $O/ncu_c5_$v.csv | cut -d, -f 5,13-16 | head -4
done
