#!/bin/bash
# A/B timing of K1 variants on a reduced C2 (N rows from $1, default 2M): pipeline-only, no-selection, full; CG 1 vs 2.
N=${1:-2000000}
B="python bench.py --n $N --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
for cg in 2 1; do for dbg in 3 1 0; do
  echo "== CG=$cg DEBUG=$dbg"
  NXF_K1_CTA_GROUP=$cg NXF_K1_DEBUG=$dbg $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('kernel_ms %.2f  TFLOP/s %.0f  frac %.3f  clocks %s' % (r['kernel_ms'], r['achieved'], r['frac'], d['clocks']))"
done; done
