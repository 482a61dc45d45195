#!/bin/bash
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
from generalsparse_b200 import synth
rows,cols,r,c=synth.c1_standin(); synth.write_mtx('/tmp/c1.mtx',rows,cols,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
run() { echo "== $*"; env $ENVV $T /tmp/c1.mtx "$@" --repeat 2000 | grep -E "kernel:|correct|wrong|c1.mtx [0-9]|min time"; }
for N in 8 16 32; do
 for G in 4 8 16; do ENVV="GS_ROW_GROUP=$G" run $N --kernel b200_row_smem; done
 ENVV="GS_ROW_GROUP=8" run $N --kernel b200_row_smem --family warp_total --sparse-coarsen 1
 ENVV="" run $N --kernel auto
 ENVV="" run $N --kernel auto --half
done
