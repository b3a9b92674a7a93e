#!/bin/bash
mkdir -p gpurun_out/r02l
O=gpurun_out/r02l
for rep in 1 2; do
  python bench.py --no-cpu --no-also --no-sustained > $O/fused_$rep.json 2> $O/fused_$rep.err
  B2NN_FUSED_UPDATE=0 python bench.py --no-cpu --no-also --no-sustained > $O/unfused_$rep.json 2> $O/unfused_$rep.err
  python bench.py --no-cpu --no-also --no-sustained --e2e-rows u8 > $O/fused_u8_$rep.json 2> $O/fused_u8_$rep.err
  B2NN_FUSED_UPDATE=0 python bench.py --no-cpu --no-also --no-sustained --e2e-rows u8 > $O/unfused_u8_$rep.json 2> $O/unfused_u8_$rep.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02l/*.json")):
    d=json.load(open(f)); print(f.split("/")[-1], "ms/step %.4f e2e %.4f"%(d["ms_per_step"], d["e2e"]["ms_per_step"]))
PY
