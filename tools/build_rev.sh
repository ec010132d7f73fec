#!/bin/bash
# Build libnexus_flat.so from a git revision into nexus_b200/libnexus_flat_<name>.so (for same-box A/B runs, tools/ab.sh)
# usage: tools/build_rev.sh <rev> <name>
set -e
REV=$1; NAME=$2; ROOT=$(cd $(dirname $0)/.. && pwd)
T=$(mktemp -d)
git -C $ROOT archive $REV nexus_b200 include | tar -x -C $T
(cd $T && python -c "
import sys; sys.path.insert(0,'$T')
from nexus_b200 import build as b
print(b.build())" | tail -1)
cp $T/nexus_b200/libnexus_flat.so $ROOT/nexus_b200/libnexus_flat_$NAME.so
rm -rf $T
echo built nexus_b200/libnexus_flat_$NAME.so from $REV
