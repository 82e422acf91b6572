#!/bin/bash
# EAM+TTM bench line with and without SM isolation of the forked lattice kernel
for iso in 1 0; do
  IMDCU_TTM_ISOLATE=$iso timeout 400 python bench.py --ttm --cells 100 100 100 --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 2>gpurun_out/err.log > gpurun_out/iso_$iso.json
  python -c "
import json;d=json.load(open('gpurun_out/iso_$iso.json'));print('iso',$iso,d['value'],d['ms_per_step'],{k:round(v['ms_total']/5,3) for k,v in d['phases'].items()})"
done
tail -3 gpurun_out/err.log
