#!/bin/bash
# round 2, 1 GPU: two energy levers of the GEMM under the power cap -- lane-0 polling of the epilogue warps (poll) and the
# banded tile order (band) -- as build variants beside the default; parity of `both`, C4 / C5 lines with the effective SM
# clock, DRAM bytes per C5 launch for default and band (ncu)
#   gpurun --timeout 1200 -- 'bash scratch/r02_call16.sh'
mkdir -p gpurun_out/r02s
O=gpurun_out/r02s
L=graph-neural-net_b200/lib
B2NN_LIB=$L/libb2nn_both.so timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_both.log 2>&1; echo "both parity rc=$?"; tail -2 $O/pytest_both.log
show() { python -c "
import json,sys; d=json.load(open('$1')); r=d['roofline']; k=r['in_kernel_timer']
print('$2: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'busy %.4f'%k['contraction_busy_ms_per_step'], 'eff MHz %.0f'%k['sm_mhz_effective'], 'frac_burst %.3f'%r['frac_burst'])"; }
for i in 1 2; do for v in def poll band both; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --no-cpu --no-also --no-sustained > $O/bench_${v}_$i.json 2> $O/bench_${v}_$i.err; show $O/bench_${v}_$i.json "c4 $v $i"
done; done
for v in def poll band both; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5_$v.json 2> $O/bench_c5_$v.err; show $O/bench_c5_$v.json "c5 $v"
done
for v in def band; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum --clock-control none -k regex:gemm_bf16 -s 20 -c 6 --csv --log-file $O/ncu_c5_$v.csv python bench.py --workload c5 --steps 2 --warmup 1 --no-cpu --no-also --no-sustained > $O/ncu_c5_$v.log 2>&1; echo "ncu $v rc=$?"
  python - <<PY
import csv
rows=[r for r in csv.reader(open('$O/ncu_c5_$v.csv')) if len(r)>10]
hdr=rows[0]; iN=hdr.index('Kernel Name'); iM=hdr.index('Metric Name'); iV=hdr.index('Metric Value'); iI=hdr.index('ID')
acc={}
for r in rows[1:]:
    acc.setdefault(r[iI],{'k':r[iN][:60]})[r[iM]]=r[iV]
for i,a in acc.items(): print('$v', i, a)
PY
done
