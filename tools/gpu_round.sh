#!/bin/bash
# One GPU-box pass: parity tests, smoke, one bench line per BASELINE configuration (five fp32 + two fp64), the ncu launch list and
# ncu --set full captures of the band kernels.  Everything lands in gpurun_out/ with the tag in the name.
# Usage (from the repo root on the box): bash tools/gpu_round.sh <tag> [quick]
TAG=${1:-r02}
QUICK=$2
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_$TAG.log
python __graft_entry__.py --smoke > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke_$TAG.log
# the metric's configuration first, with the reference-CUDA arm and the CPU arm in the same run
python bench.py > $O/bench_big_$TAG.json 2> $O/bench_big_$TAG.err; echo "bench big rc=$?"
python tools/benchline.py $O/bench_big_$TAG.json
for C in rfmip allsky rce512 reduced; do
  python bench.py --config $C --steps 5 --warmup 3 --no-ref-cuda > $O/bench_${C}_$TAG.json 2> $O/bench_${C}_$TAG.err; echo -n "$C: "; python tools/benchline.py $O/bench_${C}_$TAG.json
done
python bench.py --config rfmip --ncol 1801 --steps 5 --warmup 3 --no-ref-cuda --no-cpu > $O/bench_rfmip_onecall_$TAG.json 2>> $O/bench_rfmip_$TAG.err; echo -n "rfmip one call: "; python tools/benchline.py $O/bench_rfmip_onecall_$TAG.json
for C in big rce512; do
  python bench.py --config $C --dtype f64 --steps 3 --warmup 3 --no-cpu > $O/bench_${C}_f64_$TAG.json 2> $O/bench_${C}_f64_$TAG.err; echo -n "$C f64: "; python tools/benchline.py $O/bench_${C}_f64_$TAG.json
done
python tools/time_angles.py > $O/angles_$TAG.txt 2>&1; RB200_ANGLES=launches python tools/time_angles.py >> $O/angles_$TAG.txt 2>&1; cat $O/angles_$TAG.txt
[ -n "$QUICK" ] && exit 0
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_rfmip_$TAG.csv python bench.py --config rfmip --steps 1 --warmup 3 --no-cpu --no-e2e --no-ref-cuda > $O/ncu_list_rfmip_$TAG.log 2>&1
python tools/launch_summary.py $O/launches_rfmip_$TAG.csv > $O/launches_rfmip_${TAG}_summary.txt 2>&1
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-ref-cuda"
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/launches_$TAG.csv $B > $O/ncu_list_$TAG.log 2>&1
python tools/launch_summary.py $O/launches_$TAG.csv > $O/launches_${TAG}_summary.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:lwsw_band_kernel --launch-skip 20 -c 2 -f -o $O/prof_pair_$TAG $B > $O/ncu_pair_$TAG.log 2>&1
python tools/ncu_summary.py $O/prof_pair_$TAG.ncu-rep > $O/ncu_pair_$TAG.txt 2>&1
for K in sw lw; do
  ncu --set full --clock-control none --import-source on -k regex:${K}_band_kernel --launch-skip 20 -c 1 -f -o $O/prof_${K}_f64_$TAG $B --dtype f64 > $O/ncu_${K}_f64_$TAG.log 2>&1
  python tools/ncu_summary.py $O/prof_${K}_f64_$TAG.ncu-rep > $O/ncu_${K}_f64_$TAG.txt 2>&1
done
ls -la $O/*.ncu-rep
