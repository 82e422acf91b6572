#!/bin/bash
# Round 2, 3-D sweep variants (brick shape, CTAs per SM, loads issued up front): bit-exactness, then cell-updates/s each.
#   gpurun --timeout 1800 -- 'bash profiles/r02_call5.sh r02n'
tag=${1:-r02n}
out=gpurun_out; mkdir -p $out
for v in ${VARIANTS:-0 1 2 3 4 5 6}; do
  IMDCU_TTM3D_VARIANT=$v timeout 300 python -m pytest tests/test_ttm3d_gpu.py -m gpu -x -q > $out/${tag}_pytest_v$v.log 2>&1; rc=$?
  IMDCU_TTM3D_VARIANT=$v timeout 300 python bench.py --workload ttm3d --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_bench_v$v.json 2> $out/${tag}_bench_v$v.err
  python - <<PY
import json
try:
    d=json.loads(open("$out/${tag}_bench_v$v.json").read().strip().splitlines()[-1])
    r=d["roofline"]
    print("variant $v: pytest rc=$rc  %.4g %s  kernel %.3f ms  HBM frac %.3f" % (d["value"], d["unit"], r["ms_per_launch"], r["frac"]))
except Exception as e:
    print("variant $v: pytest rc=$rc  bench failed:", e)
PY
done
