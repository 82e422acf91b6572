#!/bin/bash
# Round 2: whole GPU suite with the new defaults (rep4 pass 1, classifier list build, sweep variant 9), the three
# bench lines, one full ncu capture of the sweep kernel.
#   gpurun --timeout 2400 -- 'bash profiles/r02_call6.sh r02o'
tag=${1:-r02o}
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q -rs > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest.log
tail -4 $out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_ttm_n1.json 2> $out/${tag}_bench_ttm_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --workload nve --steps 40 --warmup 5 --no-cpu-baseline > $out/${tag}_bench_nve_n1.json 2> $out/${tag}_bench_nve_n1.err; echo "nve rc=$?"
timeout 600 python bench.py --workload ttm3d --steps 20 --warmup 5 > $out/${tag}_bench_ttm3d.json 2> $out/${tag}_bench_ttm3d.err; echo "ttm3d rc=$?"
python - <<PY
import json
for f in ("ttm_n1","ref","nve_n1","ttm3d"):
    try:
        d=json.loads(open("$out/${tag}_bench_%s.json"%f).read().strip().splitlines()[-1])
        print(f, "%.4g %s"%(d["value"], d["unit"]), "ms/step", d.get("ms_per_step"), "e2e", d.get("e2e",{}).get("value"), "frac", (d.get("roofline") or {}).get("frac"))
    except Exception as e: print(f, "failed", e)
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_ttm3d_diff_direct' --launch-skip 10 -c 2 -f -o $out/${tag}_ttm3d_full \
  python bench.py --workload ttm3d --steps 4 --warmup 3 --e2e-steps 1 --profile-steps 1 --no-cpu-baseline > $out/${tag}_ncu_ttm3d.log 2>&1; echo "ncu rc=$?"
