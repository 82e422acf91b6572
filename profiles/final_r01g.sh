#!/bin/bash
# last pass of round 1g: whole GPU suite, launch list of the EAM+TTM step, full captures of the lattice kernel
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/r01g_pytest_final.log 2>&1; tail -2 $out/r01g_pytest_final.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/r01g_launches_ttm.csv \
  python bench.py --ttm --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/ncu_l.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ttm_grid --launch-skip 4 -c 1 -f -o $out/r01g_ttmgrid \
  python bench.py --ttm --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/ncu_g.log 2>&1; echo "ttm_grid rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ttm_grid --launch-skip 4 -c 1 -f -o $out/r01g_ttmgrid_tmm \
  python bench.py --ttm --tmm --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/ncu_t.log 2>&1; echo "ttm_grid tmm rc=$?"
