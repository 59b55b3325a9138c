#!/bin/bash
# ncu --set full of the LW and SW fused kernels (regex $1/$2), summaries only. Usage: tools/ncu_two.sh <tag> [kernel-suffix]
TAG=$1; SUF=${2:-col_kernel}; O=gpurun_out; mkdir -p $O
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --ncol 189440"
for K in sw lw; do
  ncu --set full --clock-control none --import-source on -k regex:${K}_${SUF} --launch-skip 2 -c 1 -f -o $O/prof_${K}_$TAG $B > $O/ncu_${K}_$TAG.log 2>&1
  python tools/ncu_summary.py $O/prof_${K}_$TAG.ncu-rep > $O/ncu_${K}_$TAG.txt 2>&1
  ncu -i $O/prof_${K}_$TAG.ncu-rep --page source --csv > $O/src_${K}_$TAG.csv 2>/dev/null
  rm -f $O/prof_${K}_$TAG.ncu-rep
done
