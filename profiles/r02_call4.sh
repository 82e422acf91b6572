#!/bin/bash
# Round 2, neighbour-list build experiment: k_nbl_build2 (float filter + exact test) against k_nbl_build.
#   gpurun --timeout 1800 -- 'bash profiles/r02_call4.sh r02i'
tag=${1:-r02i}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_eam_gpu.py tests/test_ttm_gpu.py tests/test_nrb_gpu.py -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest.log
tail -4 $out/${tag}_pytest.log
for f in 0 1; do
  IMDCU_NBL_FILTER=$f timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_nbl_build|k_posf' -c 8 --csv \
    --log-file $out/${tag}_nbl_f$f.csv python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_f$f.log 2>&1
  echo "filter=$f:"; grep -o '"k_nbl_build[^"]*"\|"k_posf[^"]*"\|"[0-9.]*"$' $out/${tag}_nbl_f$f.csv | paste - - | head -6
  IMDCU_NBL_FILTER=$f timeout 600 python bench.py --steps 40 --warmup 5 --e2e-steps 2 --no-cpu-baseline > $out/${tag}_bench_f$f.json 2> $out/${tag}_bench_f$f.err
  python - <<PY
import json
d=json.loads(open("$out/${tag}_bench_f$f.json").read().strip().splitlines()[-1])
print("nbl_filter=$f: %.3f ms/step, %.4g atom-steps/s, rebuilds %d in %d steps" % (d["ms_per_step"], d["value"], d["nbl_rebuilds"], d["steps"]))
PY
done
