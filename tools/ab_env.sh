#!/bin/bash
# Like ab.sh but also sweeps environment settings: tools/ab_env.sh "libs" "ENV1=a,ENV2=b ENV1=c ..." [bench args]
LIBS=$1; ENVS=$2; shift; shift
ARGS=${@:-"--steps 8 --warmup 3 --no-cpu-baseline --no-e2e"}
for e in $ENVS; do for lib in $LIBS; do
  env $(echo $e | tr ',' ' ') NEXUS_FLAT_LIB=$PWD/$lib python bench.py $ARGS 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; c=d['clocks']
print('%-24s %-34s step %.2f ms  kernel %.2f ms  %.0f TFLOP/s  sm %s MHz recall %.4f' % ('$e', '$lib', d['ms_per_step'], r['kernel_ms'], r['achieved'], c['sm_mhz'], d['check']['recall_vs_exact_fp32_kernel_64q']))"
done; done
