#!/bin/bash
# launch list of the EAM+TTM step (cold-cache, serialised: shares only)
out=gpurun_out; mkdir -p $out; tag=${1:-q}
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_ttm.csv \
  python bench.py --ttm --cells 100 100 100 --steps 3 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch_ttm.log 2>&1; echo "rc=$?"
