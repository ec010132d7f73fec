#!/bin/bash
# same-box A/B of pool kernel variants by ncu kernel durations: tools/ab_pool.sh "<libs>" <bench_pool args>
LIBS=$1; shift
for l in $LIBS; do [ "$l" = cur ] && l=""
  echo "== lib=$l $@"
  NEXUS_FLAT_LIB=$PWD/nexus_b200/libnexus_flat${l:+_$l}.so ncu --metrics gpu__time_duration.sum --kernel-name regex:pool_mean --csv python tools/bench_pool.py "$@" --iters 6 2>/dev/null | grep pool_mean | awk -F'","' '{print $NF}' | tr -d '"' | sort -n | head -12 | tr '\n' ' '; echo
done
