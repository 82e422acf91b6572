#!/bin/bash
# N-GPU pass: 2-rank parity checks, then the EAM+TTM(+NRB) bench line on N GPUs (no profiler)
#   gpurun --gpus 2 --timeout 900 -- 'bash profiles/multi_round.sh r01e 2'
tag=${1:-rXX}; n=${2:-2}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -15
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus $n --ttm --nrb --cells 100 100 100 --steps 20 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_ttm_nrb_n$n.json 2> $out/${tag}_bench_ttm_nrb_n$n.err; echo "ttm+nrb n=$n rc=$?"
python -c "import json;d=json.load(open('$out/${tag}_bench_ttm_nrb_n$n.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches'])"; tail -3 $out/${tag}_bench_ttm_nrb_n$n.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29542 \
  bench.py --gpus $n --steps 40 --warmup 5 --no-cpu-baseline > $out/${tag}_bench_n$n.json 2> $out/${tag}_bench_n$n.err; echo "nve n=$n rc=$?"
python -c "import json;d=json.load(open('$out/${tag}_bench_n$n.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches'])"
