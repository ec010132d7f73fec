#!/bin/bash
# same-box A/B of the small-batch kernels: tools/ab_small.sh "<lib names>" <sweep_small args...>
LIBS=$1; shift
for l in $LIBS; do [ "$l" = cur ] && l=""
  echo "== lib=$l"
  NEXUS_FLAT_LIB=$PWD/nexus_b200/libnexus_flat${l:+_$l}.so python tools/sweep_small.py "$@" 2>&1 | cut -c60-215
done
