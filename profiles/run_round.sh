#!/bin/bash
# One GPU-box pass: tests, bench lines (no profiler), then the ncu launch list and full captures.
#   gpurun --timeout 1500 -- 'bash profiles/run_round.sh r01b'
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
set -x
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
timeout 600 python bench.py --steps 40 --warmup 5 > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --ttm --cells 100 100 100 --steps 20 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_ttm.json 2> $out/${tag}_bench_ttm.err; echo "ttm rc=$?"
cat $out/${tag}_bench_n1.json $out/${tag}_bench_ref.json $out/${tag}_bench_ttm.json
# launch list (cold-cache, serialised): shares only
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv \
  python bench.py --steps 6 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_ttm.csv \
  python bench.py --ttm --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch_ttm.log 2>&1; echo "ncu ttm launches rc=$?"
# full captures: one launch each of the EAM passes, the integrator, the TTM kernels
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_eam_pass1|k_eam_pass2|k_move_ttm|k_ttm_grid|k_ttm_atoms1|k_ttm_atoms2' \
  --launch-skip 12 -c 6 -f -o $out/${tag}_full \
  python bench.py --ttm --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
# 3-D TTM sweep (configs[4]): bench line, then one full capture of the sweep kernel
timeout 600 python bench.py --workload ttm3d --steps 20 --warmup 3 > $out/${tag}_bench_ttm3d.json 2> $out/${tag}_bench_ttm3d.err; echo "ttm3d rc=$?"
cat $out/${tag}_bench_ttm3d.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ttm3d_diff --launch-skip 4 -c 1 -f -o $out/${tag}_ttm3d \
  python bench.py --workload ttm3d --steps 3 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_ttm3d.log 2>&1; echo "ncu ttm3d rc=$?"
ls -la $out
