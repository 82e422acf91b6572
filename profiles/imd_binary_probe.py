#!/usr/bin/env python
"""One run of IMD-with-CUDA on the bench film with IMD_CUDA_TIMING=1: where its main loop spends the time."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
r = bench.run_reference_ttm((100, 100, 100), 1, steps, 1, False, 250, binary=bench.IMD_CUDA_TTM,
                            env_extra={"IMD_CUDA_TIMING": "1", "IMD_KEEP_OUTPUT": "1"})
print({k: r[k] for k in ("steps", "main_loop_s", "wall_s", "value")}, "ms per step", 1e3 * r["main_loop_s"] / r["steps"])
print(r.get("stdout_tail", ""))
