#!/bin/bash
# last pass of the round on one B200: full GPU suite, then the bench lines that changed with the K1 short-row work
O=gpurun_out
timeout -s KILL 900 python -m pytest tests -m gpu -q 2>&1 | tail -8 > $O/r2_pytest_gpu_1gpu.log
timeout -s KILL 600 python bench.py > $O/r2_bench_c2.json 2> $O/r2_bench_c2.err
timeout -s KILL 300 python bench.py --config c5 --no-extras --no-cpu-baseline > $O/r2_bench_c5.json 2> $O/r2_bench_c5.err
timeout -s KILL 300 python bench.py --config c2f --no-extras --no-cpu-baseline > $O/r2_bench_c2f.json 2> $O/r2_bench_c2f.err
timeout -s KILL 200 python tools/sweep_small.py --config c5 --k 10 --nqs 1,8,32,64,96,128 > $O/r2_sweep_small_c5_k10.log 2>&1
timeout -s KILL 200 python tools/sweep_small.py --config c5 --nqs 1,8,32,64,96,128 > $O/r2_sweep_small_c5.log 2>&1
cat $O/r2_pytest_gpu_1gpu.log
for f in r2_bench_c2 r2_bench_c5 r2_bench_c2f; do python - <<PY
import json
d=json.loads(open('$O/$f.json').read().strip().splitlines()[-1])
print('$f', round(d['value'],1), round(d['ms_per_step'],2), round(d['e2e']['value'],1), round(d['roofline']['frac'],4), round(d['roofline']['kernel_ms'],2), d['check']['oracle_ok'], d['clocks']['sm_mhz'])
PY
done
