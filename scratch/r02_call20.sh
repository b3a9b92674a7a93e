#!/bin/bash
# round 2, 1 GPU: row-major against banded tile order on C5 at the per-GPU shape of the 8-GPU step (4096 rows) and at 32768 rows
#   gpurun --timeout 600 -- 'bash scratch/r02_call20.sh'
mkdir -p gpurun_out/r02w
O=gpurun_out/r02w
L=graph-neural-net_b200/lib
show() { python -c "
import json,sys; d=json.load(open('$1')); r=d['roofline']; k=r['in_kernel_timer']
print('$2: ms/step %.4f'%d['ms_per_step'], 'busy %.4f'%k['contraction_busy_ms_per_step'], 'eff MHz %.0f'%k['sm_mhz_effective'], 'frac_burst %.3f'%r['frac_burst'])"; }
export B2NN_BENCH_C5_ROWS=4096
for i in 1 2; do for v in rowmajor def; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --workload c5 --steps 20 --warmup 5 --no-cpu --no-also --no-sustained > $O/bench_c5r4096_${v}_$i.json 2> $O/bench_c5r4096_${v}_$i.err; show $O/bench_c5r4096_${v}_$i.json "c5 4096 rows $v $i"
done; done
export B2NN_BENCH_C5_ROWS=16384
for v in rowmajor def; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --workload c5 --steps 8 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5r16384_$v.json 2> $O/bench_c5r16384_$v.err; show $O/bench_c5r16384_$v.json "c5 16384 rows $v"
done
