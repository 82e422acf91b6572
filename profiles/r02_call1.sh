#!/bin/bash
# Round 2, first GPU pass: parity suite with the new parametrisations (observed TTM errors in the summary),
# the default bench line (EAM+TTM film, configs[2]) with its reference leg, the NVE line, launch list + full captures.
#   gpurun --timeout 2400 -- 'bash profiles/r02_call1.sh r02a'
tag=${1:-r02a}
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_gpu.txt 2>&1
nproc >> $out/${tag}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -rs --durations=15 -s > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest.log
tail -60 $out/${tag}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --workload nve --steps 40 --warmup 5 > $out/${tag}_bench_nve.json 2> $out/${tag}_bench_nve.err; echo "nve rc=$?"
cat $out/${tag}_bench_n1.json $out/${tag}_bench_ref.json $out/${tag}_bench_nve.json
tail -3 $out/${tag}_bench_n1.err $out/${tag}_bench_ref.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/${tag}_launches_ttm.csv \
  python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_eam_pass1|k_eam_pass2|k_nbl_build|k_move_ttm' \
  --launch-skip 2 -c 5 -f -o $out/${tag}_full \
  python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $out
