#!/bin/bash
# round 2, 1 GPU: epilogue look-ahead of the GEMM (B2NN_GEMM_EPI_PIPE): base = HEAD before the change (lib/libb2nn_base.so),
# default = aux loads of EPI_DX one chunk ahead, p2 = TMEM reads one chunk ahead too.  Parity of both, then the C4 line
# alternating, then the role-warp cycle counters.
#   gpurun --timeout 1200 -- 'bash scratch/r02_call13.sh'
mkdir -p gpurun_out/r02p
O=gpurun_out/r02p
L=graph-neural-net_b200/lib
timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_p1.log 2>&1; echo "p1 parity rc=$?"; tail -2 $O/pytest_p1.log
B2NN_LIB=$L/libb2nn_p2.so timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_p2.log 2>&1; echo "p2 parity rc=$?"; tail -2 $O/pytest_p2.log
for i in 1 2 3; do for v in base p1 p2; do
  if [ $v = p1 ]; then unset B2NN_LIB; else export B2NN_LIB=$L/libb2nn_$v.so; fi
  python bench.py --no-cpu --no-also --no-sustained > $O/bench_${v}_$i.json 2> $O/bench_${v}_$i.err
  python -c "
import json; d=json.load(open('$O/bench_${v}_$i.json')); r=d['roofline']
print('$v run $i: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'u8 %.4f'%d['e2e']['pixel_byte_rows']['ms_per_step'], 'avg launch per-run %.4f'%r['avg_launch_ms'], 'frac_burst %.3f'%r['frac_burst'], 'in-kernel busy %.4f'%r['in_kernel_timer']['contraction_busy_ms_per_step'])"
done; done
unset B2NN_LIB
for v in base p1 p2; do
  if [ $v = p1 ]; then unset B2NN_LIB; else export B2NN_LIB=$L/libb2nn_$v.so; fi
  B2NN_GEMM_STALLS=1 python scratch/stalls.py 2>&1 | tail -8 | cut -c1-330 > $O/stalls_$v.log; echo "== $v"; cat $O/stalls_$v.log
done
unset B2NN_LIB
B2NN_TRACE=1 python bench.py --steps 8 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_trace.json 2> $O/bench_trace.err; grep -c . $O/bench_trace.err
