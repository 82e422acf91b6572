#!/bin/bash
# time k_ttm_grid for every library variant under imd_b200/variants (built by hand from git revisions)
cp imd_b200/libimdcu.so /tmp/libimdcu_keep.so
for f in imd_b200/variants/libimdcu_*.so; do
  v=$(basename $f .so); v=${v#libimdcu_}
  cp $f imd_b200/libimdcu.so
  echo "$v $(PYTHONPATH=. timeout 300 python profiles/ttm_grid_probe.py 100 250 40 2>&1 | tail -1)"
done
cp /tmp/libimdcu_keep.so imd_b200/libimdcu.so
