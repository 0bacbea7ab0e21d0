#!/bin/bash
# A/B of scan kernel shapes/variants (dev only)
cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu > gpurun_out/t.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/t.log
tail -3 gpurun_out/t.log
for sh in 0 1; do for v in 0 1 2 3; do
  VSV_SCAN_SHAPE=$sh VSV_SCAN_VAR=$v python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ab_${sh}_${v}.log 2>&1
  python - <<PY
import json
for l in open("gpurun_out/ab_${sh}_${v}.log"):
    if l.startswith("{"):
        j=json.loads(l); print("shape",$sh,"var",$v,"step_ms",round(j["ms_per_step"],3),"scan_ms",round(j["roofline"]["kernel_ms"],4),"frac",round(j["roofline"]["frac"],4),"e2e_ms",round(j["e2e"]["ms_per_step"],2), "launches", j["gpu_launches"])
PY
done; done
