#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench, ncu launch list, ncu --set full of the two band kernels.
# Usage (from the repo root on the box): bash tools/gpu_round.sh <tag>
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_$TAG.log
python __graft_entry__.py --smoke > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke_$TAG.log
python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench rc=$?"
tail -c 600 $O/bench_$TAG.json
if [ -s $O/bench_$TAG.json ]; then
  B="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e"
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$TAG.csv $B > $O/ncu_list_$TAG.log 2>&1
  for K in sw lw; do
    ncu --set full --clock-control none --import-source on -k regex:${K}_band_kernel --launch-skip 30 -c 2 -f -o $O/prof_${K}_$TAG $B > $O/ncu_${K}_$TAG.log 2>&1
    python tools/ncu_summary.py $O/prof_${K}_$TAG.ncu-rep > $O/ncu_${K}_$TAG.txt 2>&1
  done
  ls -la $O/*.ncu-rep
fi
