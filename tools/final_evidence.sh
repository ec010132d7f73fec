#!/bin/bash
# Round-end evidence on one B200: bench lines (C2, C2 fp32 storage, C5, reference arm), small-batch sweeps, ncu launch list.
# Every step is bounded; outputs land in gpurun_out/ and are copied to profiles/ by hand.
O=gpurun_out
timeout -s KILL 600 python bench.py > $O/r2_bench_c2.json 2> $O/r2_bench_c2.err
timeout -s KILL 400 python bench.py --impl reference > $O/r2_bench_c2_reference_arm.json 2> $O/r2_bench_ref.err
timeout -s KILL 400 python bench.py --config c2f --no-extras --no-cpu-baseline > $O/r2_bench_c2f.json 2> $O/r2_bench_c2f.err
timeout -s KILL 300 python bench.py --config c5 --no-extras --no-cpu-baseline > $O/r2_bench_c5.json 2> $O/r2_bench_c5.err
timeout -s KILL 200 python tools/sweep_small.py --config c2 > $O/r2_sweep_small_c2.log 2>&1
timeout -s KILL 200 python tools/sweep_small.py --config c2 --store f32 > $O/r2_sweep_small_c2_f32.log 2>&1
timeout -s KILL 200 python tools/sweep_small.py --config c5 --nqs 1,8,32,64,96,128 > $O/r2_sweep_small_c5.log 2>&1
timeout -s KILL 200 python tools/sweep_small.py --config c5 --k 10 --nqs 1,8,32,64,96,128 > $O/r2_sweep_small_c5_k10.log 2>&1
timeout -s KILL 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_ncu_launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > $O/r2_ncu_run.log 2>&1
for f in r2_bench_c2 r2_bench_c2_reference_arm r2_bench_c2f r2_bench_c5; do echo "== $f"; cut -c1-400 $O/$f.json; done
tail -2 $O/r2_sweep_small_c5_k10.log | cut -c1-300
