#!/bin/bash
# round 2, 1 GPU: banded tile order (default now) with / without the L2 eviction hints on large launches: parity of the
# default build, C5 and C4 lines, DRAM bytes per C5 launch (ncu)
#   gpurun --timeout 1200 -- 'bash scratch/r02_call17.sh'
mkdir -p gpurun_out/r02t
O=gpurun_out/r02t
L=graph-neural-net_b200/lib
timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_default.log 2>&1; echo "default parity rc=$?"; tail -2 $O/pytest_default.log
show() { python -c "
import json,sys; d=json.load(open('$1')); r=d['roofline']; k=r['in_kernel_timer']
print('$2: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'busy %.4f'%k['contraction_busy_ms_per_step'], 'eff MHz %.0f'%k['sm_mhz_effective'], 'frac_burst %.3f'%r['frac_burst'])"; }
for i in 1 2; do for v in nohint def; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5_${v}_$i.json 2> $O/bench_c5_${v}_$i.err; show $O/bench_c5_${v}_$i.json "c5 $v $i"
done; done
for v in nohint def; do
  unset B2NN_LIB; if [ $v != def ]; then export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --no-cpu --no-also --no-sustained > $O/bench_c4_$v.json 2> $O/bench_c4_$v.err; show $O/bench_c4_$v.json "c4 $v"
done
unset B2NN_LIB
python bench.py --precision tf32 --no-cpu --no-also --no-sustained > $O/bench_tf32.json 2> $O/bench_tf32.err; show $O/bench_tf32.json "c4 tf32 def"
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum --clock-control none -k regex:gemm_bf16 -s 20 -c 17 --csv --log-file $O/ncu_c5_def.csv python bench.py --workload c5 --steps 2 --warmup 1 --no-cpu --no-also --no-sustained > $O/ncu_c5_def.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open('$O/ncu_c5_def.csv')) if len(r)>10]
hdr=rows[0]; iN=hdr.index('Kernel Name'); iM=hdr.index('Metric Name'); iV=hdr.index('Metric Value'); iI=hdr.index('ID')
acc={}
for r in rows[1:]:
    acc.setdefault(r[iI],{'k':r[iN][26:36]})[r[iM].split('__')[0]+r[iM][-9:]]=r[iV]
for i,a in acc.items(): print(i, a)
PY
