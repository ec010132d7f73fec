#!/bin/bash
# K2T on the C5 corpus (10M x 64 bf16, 128-byte rows): per-launch durations (ncu, gpu__time_duration) of the shipped library and of
# the knock-out builds of search_stream_tc.cu, to see which stage paces the stream.  Run on the GPU box from the repo root:
#   for n in 1 2 3 5; do tools/build_variant.sh k2t_ko$n "-DNXF_K2T_KNOCKOUT=$n"; done     (here, before gpurun)
#   bash tools/k2t_knockout.sh [nq] [k]
# Every run is bounded by `timeout -s KILL`: a knock-out build computes nothing valid and a hung kernel must not hold the box.
# Output: one line per configuration = the durations (ns) of the k2t_stream_kernel launches of two searches
# (estimate, sample, main, fall-back; the fall-back is a no-op unless the counts came back short, as they do in builds 1-3 and 5).
NQS=${1:-128}; K=${2:-10}
L=gpurun_out/k2t_knockout_nq${NQS}_k${K}.log; rm -f $L
R=$(cd $(dirname $0)/.. && pwd)/nexus_b200
run() { tag=$1; shift; echo "== $tag: $*" >> $L; env "$@" timeout -s KILL 100 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k2t_stream -c 8 --csv python -u tools/sweep_small.py --config c5 --k $K --nqs $NQS --iters 2 2>&1 | grep -v "^==" | awk -F'","' '/k2t_stream/ {print $(NF)}' | sed 's/"//g' | tr '\n' ' ' >> $L; echo >> $L; }
run shipped NXF_K2T_FOLD=0
run fold NXF_K2T_FOLD=1
for n in 2 1 3 5; do
  [ -f $R/libnexus_flat_k2t_ko$n.so ] && run ko$n NEXUS_FLAT_LIB=$R/libnexus_flat_k2t_ko$n.so
done
[ -f $R/libnexus_flat_k2t_ko5.so ] && run ko5_fold NEXUS_FLAT_LIB=$R/libnexus_flat_k2t_ko5.so NXF_K2T_FOLD=1
cat $L
