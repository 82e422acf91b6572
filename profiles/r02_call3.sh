#!/bin/bash
# Round 2, EAM pass-1 experiment: k_eam_pass1_rep4 (raw table rows, four bank-staggered copies) against
# k_eam_pass1_fused (48-byte rows).  Parity first, then the bench line per variant, then one full ncu capture.
#   gpurun --timeout 2400 -- 'bash profiles/r02_call3.sh r02f'
tag=${1:-r02f}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_eam_gpu.py -m gpu -x -q -s > $out/${tag}_pytest_eam.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_eam.log
grep -n "rep4\|passed\|failed" $out/${tag}_pytest_eam.log | tail -6
for v in "1 896" "2 896" "2 768"; do
  set -- $v
  IMDCU_FUSED_TABLE=$1 IMDCU_REP4_BLOCK=$2 timeout 600 python bench.py --steps 20 --warmup 5 --e2e-steps 2 --no-cpu-baseline \
    > $out/${tag}_bench_f$1_b$2.json 2> $out/${tag}_bench_f$1_b$2.err; echo "bench fused=$1 block=$2 rc=$?"
  python - <<PY
import json
d=json.loads(open("$out/${tag}_bench_f$1_b$2.json").read().strip().splitlines()[-1])
ph=d["phases"]
print("fused=$1 block=$2: %.3f ms/step, %.4g atom-steps/s, pass1 %.3f ms, pass2 %.3f ms (profiled)" % (d["ms_per_step"], d["value"], ph["eam_pass1"]["ms_total"]/ph["eam_pass1"]["launches"], ph["eam_pass2"]["ms_total"]/ph["eam_pass2"]["launches"]))
PY
done
IMDCU_FUSED_TABLE=2 IMDCU_REP4_BLOCK=768 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_eam_pass1' \
  --launch-skip 2 -c 2 -f -o $out/${tag}_rep4_full \
  python bench.py --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $out | tail -12
