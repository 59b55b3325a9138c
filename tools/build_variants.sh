#!/bin/bash
# Build experiment variants of the library side by side: tools/build_variants.sh name1:"-DX -DY" name2:"-DZ" ...
# -> rte_rrtmgp_cpp_b200/librrtmgp_b200_<name>.so (git-ignored, travels to the GPU box); select with RB200_LIB.
for v in "$@"; do
  name=${v%%:*}; defs=${v#*:}
  RB200_TAG=_$name RB200_DEFS="$defs" RB200_LIB=$PWD/rte_rrtmgp_cpp_b200/librrtmgp_b200_$name.so python -m rte_rrtmgp_cpp_b200.build --force -v 2> /tmp/ptxas_$name.log > /dev/null &
done
wait
for v in "$@"; do name=${v%%:*}; echo "== $name"; grep -A1 "band_kernelIfLi4ELi16ELb0ELi9E" /tmp/ptxas_$name.log | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores" | sort | uniq -c | head -4; done
