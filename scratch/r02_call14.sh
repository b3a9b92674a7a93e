#!/bin/bash
# round 2, 1 GPU: independent contraction launches start under the previous kernel's tail (gemm_dep / B2NN_PDL_CHAIN).
# Parity with it on (bit-equality test included), then the C4 line: base = HEAD before the epilogue look-ahead and the chain
# (lib/libb2nn_base.so), chain0 = this build with B2NN_PDL_CHAIN=0, chain1 = this build (default); alternating.
#   gpurun --timeout 1200 -- 'bash scratch/r02_call14.sh'
mkdir -p gpurun_out/r02q
O=gpurun_out/r02q
L=graph-neural-net_b200/lib
timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_chain.log 2>&1; echo "chain parity rc=$?"; tail -3 $O/pytest_chain.log
for i in 1 2 3; do for v in base chain0 chain1; do
  unset B2NN_LIB B2NN_PDL_CHAIN
  if [ $v = base ]; then export B2NN_LIB=$L/libb2nn_base.so; fi
  if [ $v = chain0 ]; then export B2NN_PDL_CHAIN=0; fi
  python bench.py --no-cpu --no-also --no-sustained > $O/bench_${v}_$i.json 2> $O/bench_${v}_$i.err
  python -c "
import json; d=json.load(open('$O/bench_${v}_$i.json')); r=d['roofline']
print('$v run $i: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'u8 %.4f'%d['e2e']['pixel_byte_rows']['ms_per_step'], 'avg launch per-run %.4f'%r['avg_launch_ms'], 'frac_burst %.3f'%r['frac_burst'], 'in-kernel busy %.4f step %.4f'%(r['in_kernel_timer']['contraction_busy_ms_per_step'], r['in_kernel_timer']['step_ms']))"
done; done
unset B2NN_LIB B2NN_PDL_CHAIN
for v in chain0 chain1; do
  unset B2NN_PDL_CHAIN; if [ $v = chain0 ]; then export B2NN_PDL_CHAIN=0; fi
  python bench.py --precision tf32 --no-cpu --no-also --no-sustained > $O/bench_tf32_$v.json 2> $O/bench_tf32_$v.err; python -c "
import json; d=json.load(open('$O/bench_tf32_$v.json')); print('tf32 $v ms/step %.4f'%d['ms_per_step'])"
  python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5_$v.json 2> $O/bench_c5_$v.err; python -c "
import json; d=json.load(open('$O/bench_c5_$v.json')); print('c5 $v ms/step %.4f'%d['ms_per_step'])"
done
