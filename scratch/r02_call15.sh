#!/bin/bash
# round 2, 1 GPU: the SM clock the timed region really runs at (in-kernel cycles / ns), for 20 / 50 / 200 / 800-step regions,
# and the role-warp cycle counters of ISOLATED launches beside it
#   gpurun --timeout 900 -- 'bash scratch/r02_call15.sh'
mkdir -p gpurun_out/r02r
O=gpurun_out/r02r
timeout 300 python -m pytest tests/test_gpu_wide.py -x -q -m gpu -k "stamps or profile" > $O/pytest_stamps.log 2>&1; echo "stamps test rc=$?"; tail -2 $O/pytest_stamps.log
for s in 20 50 200 800; do
  python bench.py --steps $s --no-cpu --no-also --no-sustained > $O/bench_s$s.json 2> $O/bench_s$s.err
  python -c "
import json; d=json.load(open('$O/bench_s$s.json')); r=d['roofline']; k=r['in_kernel_timer']
print('steps $s: ms/step %.4f'%d['ms_per_step'], 'e2e %.4f'%d['e2e']['ms_per_step'], 'stamped step %.4f busy %.4f avg launch %.4f'%(k['step_ms'], k['contraction_busy_ms_per_step'], k['avg_launch_ms']), 'effective MHz %.0f'%k['sm_mhz_effective'], 'smi', d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu --no-also --no-sustained > $O/bench_c5.json 2> $O/bench_c5.err; python -c "
import json; d=json.load(open('$O/bench_c5.json')); k=d['roofline']['in_kernel_timer']; print('c5 ms/step %.3f'%d['ms_per_step'], 'effective MHz %.0f'%k['sm_mhz_effective'], 'frac_burst %.3f'%d['roofline']['frac_burst'])"
python bench.py --precision tf32 --no-cpu --no-also --no-sustained > $O/bench_tf32.json 2> $O/bench_tf32.err; python -c "
import json; d=json.load(open('$O/bench_tf32.json')); k=d['roofline']['in_kernel_timer']; print('tf32 ms/step %.3f'%d['ms_per_step'], 'effective MHz %.0f'%k['sm_mhz_effective'])"
