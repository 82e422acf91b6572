#!/bin/bash
# per-phase times of the NVE bench for every library variant under imd_b200/variants
#   gpurun -- 'bash profiles/eam_variants_run.sh > gpurun_out/eam_variants.txt'
cp imd_b200/libimdcu.so /tmp/libimdcu_keep.so
for f in /tmp/libimdcu_keep.so imd_b200/variants/libimdcu_*.so; do
  v=$(basename $f .so); v=${v#libimdcu_}
  [ $f != /tmp/libimdcu_keep.so ] && cp $f imd_b200/libimdcu.so
  timeout 300 python bench.py --steps 20 --warmup 5 --e2e-steps 1 --profile-steps 5 --no-cpu-baseline $EXTRA 2>/tmp/err.txt | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
ph = d['phases']
print('$v', 'ms/step %.3f' % d['ms_per_step'], ' '.join('%s %.3f' % (k, v['ms_total'] / 5) for k, v in ph.items()))
" || tail -3 /tmp/err.txt
done
cp /tmp/libimdcu_keep.so imd_b200/libimdcu.so
