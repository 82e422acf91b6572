#!/bin/bash
# Round 2, last profiling pass on the final tree: launch list of the default bench command (shares of the step) and
# one full capture each of the three kernels that carry the step and the rebuild.
#   gpurun --timeout 1500 -- 'bash profiles/r03_call.sh r03g'
tag=${1:-r03g}
out=gpurun_out; mkdir -p $out
timeout 600 python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_plain.json 2> $out/${tag}_plain.err; echo "plain rc=$?"
timeout 700 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $out/${tag}_launches_ttm.csv \
  python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 700 ncu --set full --clock-control none --import-source on -k regex:'k_eam_pass1_rep4|k_eam_pass2|k_nbl_build2' \
  --launch-skip 3 -c 4 -f -o $out/${tag}_full \
  python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $out | tail -8
