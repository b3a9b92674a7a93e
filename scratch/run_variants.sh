#!/bin/bash
# usage: run_variants.sh <tag>   -- 1-GPU C4 bench under the scheduling knobs
for cfg in "0 1" "1 1" "0 0" "1 0"; do set -- $cfg
  B2NN_DYNAMIC_TILES=$1 B2NN_BACKWARD_REORDER=$2 timeout 300 python bench.py --steps 50 --warmup 10 --no-cpu --no-also > gpurun_out/m_c4_d$1_r$2.json 2> gpurun_out/m_c4_d$1_r$2.err
done
