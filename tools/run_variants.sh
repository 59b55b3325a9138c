#!/bin/bash
# Bench experiment variants built by tools/build_variants.sh on the GPU box: tools/run_variants.sh base name1 name2 ...
# ("base" = the default library).  One line per variant in gpurun_out/variants.log.
O=gpurun_out; mkdir -p $O
ARGS=${BENCH_ARGS:---steps 3 --warmup 2 --no-cpu --no-e2e --no-ref-cuda}
for name in "$@"; do
  if [ "$name" = base ]; then unset RB200_LIB; else export RB200_LIB=$PWD/rte_rrtmgp_cpp_b200/librrtmgp_b200_$name.so; fi
  echo -n "$name: " | tee -a $O/variants.log
  python bench.py $ARGS 2> $O/variant_$name.err | python tools/benchline.py | tee -a $O/variants.log
done
