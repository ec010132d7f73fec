#!/bin/bash
# Build the working tree with extra nvcc flags into nexus_b200/libnexus_flat_<name>.so (same-box A/B of compile-time variants)
# usage: tools/build_variant.sh <name> "<nvcc flags>"
set -e
NAME=$1; FLAGS=$2; ROOT=$(cd $(dirname $0)/.. && pwd)
T=$(mktemp -d)
mkdir -p $T/nexus_b200 && cp -r $ROOT/nexus_b200/csrc $ROOT/nexus_b200/*.py $T/nexus_b200/ && cp -r $ROOT/include $T/
(cd $T && NXF_EXTRA_NVCC_FLAGS="$FLAGS" python -c "
import sys; sys.path.insert(0,'$T')
from nexus_b200 import build as b
print(b.build(force=True))" | tail -1)
cp $T/nexus_b200/libnexus_flat.so $ROOT/nexus_b200/libnexus_flat_$NAME.so
rm -rf $T
echo built nexus_b200/libnexus_flat_$NAME.so with "$FLAGS"
