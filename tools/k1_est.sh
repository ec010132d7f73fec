#!/bin/bash
# K1 threshold estimation experiment: seeded passes with the guaranteed bound (k'-th of the sample) vs the sampled rank
N=${1:-8800000}; shift
for kk in ${@:-100}; do for cfg in "NXF_K1_SEED=0" "NXF_K1_SEED=1" "NXF_K1_SEED=1 NXF_K1_EST=1"; do
  echo "== n=$N k=$kk $cfg"
  env $cfg python bench.py --n $N --k $kk --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms_per_step', round(d['ms_per_step'],3), 'kernel_ms(sum/launch)', round(d['roofline']['kernel_ms'],3), d['roofline']['kernel_launches_timed'], 'check', d['check'])"
done; done
