#!/bin/bash
# round 2, 1 GPU: (1) the two roofline-window modes, (2) parity of the tensor-core path with programmatic dependent launch
# on (B2NN_PDL=1), (3) A/B of the C4 bench line with it off / on, alternating on the same box.
#   gpurun --timeout 900 -- 'bash scratch/r02_pdl.sh'
mkdir -p gpurun_out/r02pdl
O=gpurun_out/r02pdl
python -m pytest tests/test_gpu_wide.py -x -q -m gpu -k "profile_windows" > $O/pytest_profile.log 2>&1; echo "profile test rc=$?"; tail -2 $O/pytest_profile.log
B2NN_PDL=1 timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_z_tf32.py tests/test_gpu_zz_fullsize.py -x -q -m gpu > $O/pytest_pdl.log 2>&1; echo "pdl parity rc=$?"; tail -3 $O/pytest_pdl.log
for i in 1 2; do for p in 0 1; do
  B2NN_PDL=$p python bench.py --no-cpu --no-also --no-sustained > $O/bench_pdl${p}_$i.json 2> $O/bench_pdl${p}_$i.err
  python -c "
import json; d=json.load(open('$O/bench_pdl${p}_$i.json')); r=d['roofline']
print('PDL=$p run $i: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'u8 %.4f'%d['e2e']['pixel_byte_rows']['ms_per_step'], 'avg launch grouped %.4f per-launch %.4f'%(r['avg_launch_ms'], r['avg_launch_ms_event_per_launch']), 'frac_burst %.3f'%r['frac_burst'], 'serialised %.4f / %.4f'%(r['serialised_step_ms'], r['serialised_step_ms_event_per_launch']))"
done; done
B2NN_PDL=1 python bench.py --precision tf32 --no-cpu --no-also --no-sustained > $O/bench_tf32_pdl1.json 2> $O/bench_tf32_pdl1.err
B2NN_PDL=0 python bench.py --precision tf32 --no-cpu --no-also --no-sustained > $O/bench_tf32_pdl0.json 2> $O/bench_tf32_pdl0.err
python -c "
import json
for p in (0,1):
    d=json.load(open('$O/bench_tf32_pdl%d.json'%p)); print('tf32 PDL=%d ms/step %.4f frac_burst %.3f'%(p, d['ms_per_step'], d['roofline']['frac_burst']))"
