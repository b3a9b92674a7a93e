#!/bin/bash
# round 2, 1 GPU: why does the BF16-row end-to-end leg run at 1.2-2.4 ms/step in ~1 of 8 bench runs (the pixel-byte leg
# never does)?  Host topology as the guest sees it, then the leg alone, 10 times free-running and under CPU masks, with the
# two streams' timeline.
#   gpurun --timeout 900 -- 'bash scratch/r02_call18.sh'
mkdir -p gpurun_out/r02u
O=gpurun_out/r02u
( nvidia-smi topo -m; lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)"; for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/vendor 2>/dev/null)" = "0x10de" ]; then echo "$d numa_node=$(cat $d/numa_node) class=$(cat $d/class)"; fi; done; ls /sys/devices/system/node/; cat /sys/devices/system/node/node*/cpulist; nproc; taskset -p $$; nvidia-smi -q | grep -i -A3 "GPU Link Info" | head -12 ) > $O/topology.txt 2>&1
head -40 $O/topology.txt
for i in 1 2 3 4 5 6 7 8 9 10; do B2NN_TRACE=1 python scratch/e2e_numa.py free$i 2> $O/trace_free$i.err | tail -1; done
N=$(nproc)
for m in "0-$((N/4-1))" "$((N/2))-$((N/2+N/4-1))" "$((N-N/4))-$((N-1))"; do for i in 1 2 3; do B2NN_TRACE=1 taskset -c $m python scratch/e2e_numa.py mask$m 2> $O/trace_mask${m}_$i.err | tail -1; done; done
grep -l . $O/*.err | head -3
