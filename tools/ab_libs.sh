#!/bin/bash
# A/B of library builds on the metric's configuration: ms per step, band kernels, everything else (prologues, zeroing, net flux).
# Usage (GPU box): bash tools/ab_libs.sh librrtmgp_b200.so librrtmgp_b200_<tag>.so ...   [extra bench.py flags in $AB_FLAGS]
for L in "$@"; do
  RB200_LIB=/root/repo/rte_rrtmgp_cpp_b200/$L python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-ref-cuda $AB_FLAGS > gpurun_out/ab_$L.json 2>/dev/null
  python - "$L" <<PY
import json, sys
d = json.loads(open("gpurun_out/ab_" + sys.argv[1] + ".json").read().strip().splitlines()[-1])
b = d["breakdown_ms_per_step"]
print(f"{sys.argv[1]:36s} step {d['ms_per_step']:8.2f} ms  band {b['band_kernels']:8.2f}  other {b['other']:7.2f}")
PY
done
