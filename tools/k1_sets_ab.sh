#!/bin/bash
# K1 with one vs two epilogue warp sets (NXF_K1_SETS) on the same box, per row length: where does the second set stop paying?
# usage: tools/k1_sets_ab.sh "128 256 384 448 512" [config] [extra bench args]
DIMS=${1:-"128 256 384 448 512"}
CFG=${2:-c2}
shift; shift
for d in $DIMS; do for s in 1 2; do
  NXF_K1_SETS=$s python bench.py --config $CFG --dim $d --n 4000000 --nq 16384 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e $@ 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('dim %4d sets %d  step %.2f ms  %.0f TFLOP/s  qps %.0f  recall %.4f' % ($d, $s, d['ms_per_step'], r['achieved'], d['value'], d['check']['recall_vs_exact_fp32_kernel_64q']))"
done; done
