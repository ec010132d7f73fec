#!/bin/bash
# prune / end-merge cycle counts of K1 (NXF_K1_DEBUG bit 5) on a shard of C2
N=${1:-1100000}; shift
for kk in ${@:-100}; do for dbg in 32 96; do
  echo "== n=$N k=$kk debug=$dbg (bit 6 = no prefetch)"
  NXF_K1_DEBUG=$dbg python bench.py --n $N --k $kk --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep -E "nxf K1 prof" | tail -1
done; done
