#!/bin/bash
# A/B timing of K1 variants on C2 with N rows ($1, default 2M): NXF_K1_DEBUG bits (1 no selection, 2 no TMEM loads,
# 4 no TMA loads, 8 no MMAs) x cta_group.  usage: tools/k1_variants.sh [N] ["dbg list"] ["cg list"]
N=${1:-2000000}
DBGS=${2:-"3 1 0"}
CGS=${3:-"2 1"}
B="python bench.py --n $N --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
for cg in $CGS; do for dbg in $DBGS; do
  echo "== CG=$cg DEBUG=$dbg"
  NXF_K1_CTA_GROUP=$cg NXF_K1_DEBUG=$dbg $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('kernel_ms %.2f  TFLOP/s %.0f  frac %.3f  clocks %s' % (r['kernel_ms'], r['achieved'], r['frac'], d['clocks']))"
done; done
