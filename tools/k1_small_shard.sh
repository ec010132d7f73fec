#!/bin/bash
# K1 on a 1/8 shard of C2 (what each rank of an 8-GPU search runs): where does the N-independent part of the kernel time go?
N=${1:-1100000}
run() { echo "== $*"; env "$@" python bench.py --n $N --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms_per_step', round(d['ms_per_step'],3), 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'sm_mhz', d['clocks']['sm_mhz'])"; }
run NXF_K1_DEBUG=0
run NXF_K1_DEBUG=16
run NXF_K1_DEBUG=1
run NXF_K1_CAP_MULT=4
run NXF_K1_CAP_MULT=8
run NXF_K1_SEED=1
run NXF_K1_ROUND_MB=16
run NXF_K1_ROUND_MB=64
