#!/bin/bash
# quick TTM pass: parity tests of the 1-D TTM, the EAM+TTM bench line (no profiler), then the launch list
out=gpurun_out; mkdir -p $out; tag=${1:-q}
timeout 600 python -m pytest tests/test_ttm_gpu.py -x -q 2>&1 | tail -5
timeout 300 python bench.py --ttm --cells 100 100 100 --steps 20 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_ttm.json 2> $out/${tag}_bench_ttm.err; echo "ttm rc=$?"
python -c "import json;d=json.load(open('$out/${tag}_bench_ttm.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],{k:v['ms_total'] for k,v in d['phases'].items()})"; tail -5 $out/${tag}_bench_ttm.err
bash profiles/quick_ttm_launches.sh $tag
