#!/bin/bash
# round 2, 1 GPU: (1) the allocation-free entry point's test alone in a fresh process, (2) the whole GPU suite (that test
# last), (3) the second version of the programmatic dependent launch (released when the producer is done): parity with it
# on, then the C4 bench line off / on, alternating.
#   gpurun --timeout 1200 -- 'bash scratch/r02_noalloc_pdl2.sh'
mkdir -p gpurun_out/r02na
O=gpurun_out/r02na
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/pytest_noalloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -4 $O/pytest_noalloc_alone.log
python -m pytest tests -x -q -m gpu > $O/pytest_gpu_all.log 2>&1; echo "gpu suite rc=$?"; tail -4 $O/pytest_gpu_all.log
B2NN_PDL=1 timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_pdl.log 2>&1; echo "pdl parity rc=$?"; tail -2 $O/pytest_pdl.log
for i in 1 2 3; do for p in 0 1; do
  B2NN_PDL=$p python bench.py --no-cpu --no-also --no-sustained > $O/bench_pdl${p}_$i.json 2> $O/bench_pdl${p}_$i.err
  python -c "
import json; d=json.load(open('$O/bench_pdl${p}_$i.json')); r=d['roofline']
print('PDL=$p run $i: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'u8 %.4f'%d['e2e']['pixel_byte_rows']['ms_per_step'], 'avg launch per-run %.4f per-launch %.4f'%(r['avg_launch_ms'], r['avg_launch_ms_event_per_launch']), 'frac_burst %.3f'%r['frac_burst'])"
done; done
