#!/bin/bash
# round 2, last GPU seconds: C5 with every second wave of tiles walking K downwards (B2NN_GEMM_KSNAKE=1 build variant)
mkdir -p gpurun_out/r02x
O=gpurun_out/r02x
L=graph-neural-net_b200/lib
for v in def ksnake; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  timeout 40 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5_$v.json 2> $O/bench_c5_$v.err
  python -c "
import json; d=json.load(open('$O/bench_c5_$v.json')); k=d['roofline']['in_kernel_timer']; print('c5 $v: ms/step %.3f'%d['ms_per_step'], 'eff MHz %.0f'%k['sm_mhz_effective'], 'last loss', d['e2e'].get('last_loss'))"
done
