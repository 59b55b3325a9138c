#!/bin/bash
# Static SASS opcode histogram of the fp32 band kernels (no GPU needed). Usage: tools/sass_hist.sh <tag>
TAG=${1:-r02}
for k in lw sw; do
  f=rte_rrtmgp_cpp_b200/csrc/_obj/fused_$k.o
  fn="_ZN5rb20014${k}_band_kernelIfLi4ELi16ELb0ELi9EEEvNS_7DevPlanIT_EENS_10BandLaunchIS2_EENS_9TileStateIS2_EENS_9FusedArgsIS2_EE"
  { echo "${k}_band_kernel<float, 4, 16, false, 9>  (cuobjdump -sass, static instruction counts by opcode)";
    cuobjdump -sass $f -fun "$fn" | grep -E "^\s+/\*[0-9a-f]{4}\*/" | awk '{for(i=2;i<=NF;i++) if ($i !~ /^@/) {print $i; break}}' | sed 's/\..*//;s/;//' | sort | uniq -c | sort -rn; } > profiles/sass_${k}_band_$TAG.txt
done
