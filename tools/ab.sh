#!/bin/bash
# A/B timing of two builds of libnexus_flat.so ON THE SAME BOX (boxes differ in sustained clocks, so numbers from different
# gpurun calls are not comparable).  usage: tools/ab.sh "libA.so libB.so ..." [rounds] [bench args...]
LIBS=${1:-"nexus_b200/libnexus_flat_base.so nexus_b200/libnexus_flat.so"}
ROUNDS=${2:-2}
shift; shift
ARGS=${@:-"--steps 8 --warmup 3 --no-cpu-baseline --no-e2e"}
for r in $(seq $ROUNDS); do for lib in $LIBS; do
  NEXUS_FLAT_LIB=$PWD/$lib python bench.py $ARGS 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; c=d['clocks']
print('%-40s step %.2f ms  kernel %.2f ms  %.0f TFLOP/s  qps %.0f  sm %s MHz %s recall %.4f' % ('$lib', d['ms_per_step'], r['kernel_ms'], r['achieved'], d['value'], c['sm_mhz'], c['reasons'], d['check']['recall_vs_exact_fp32_kernel_64q']))"
done; done
