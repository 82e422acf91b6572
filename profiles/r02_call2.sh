#!/bin/bash
# Round 2, second GPU pass: the whole parity suite including IMD-with-CUDA-hot-path (tests/test_integration_gpu.py:
# .eng and .chkpt of integration/_bin/imd_eam_nve_nbl_cuda against the unmodified reference, restart of the plain
# reference from the CUDA build's checkpoint), smoke(), the default bench line (EAM+TTM film, configs[2]) + reference arm.
#   gpurun --timeout 2400 -- 'bash profiles/r02_call2.sh r02b'
tag=${1:-r02b}
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_gpu.txt 2>&1
nproc >> $out/${tag}_gpu.txt
timeout 300 python -m pytest tests/test_integration_gpu.py -m gpu -q -rs -s > $out/${tag}_integration.log 2>&1; echo "integration rc=$?" >> $out/${tag}_integration.log
tail -25 $out/${tag}_integration.log
timeout 1500 python -m pytest tests -m gpu -x -q -rs --durations=10 -s > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest.log
tail -30 $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
tail -5 $out/${tag}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
cat $out/${tag}_bench_n1.json $out/${tag}_bench_ref.json
tail -n 3 $out/${tag}_bench_n1.err $out/${tag}_bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/${tag}_launches_ttm.csv \
  python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
ls -la $out
