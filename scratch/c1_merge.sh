#!/bin/bash
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
from generalsparse_b200 import synth
rows,cols,r,c=synth.c1_standin(); synth.write_mtx('/tmp/c1.mtx',rows,cols,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
run() { echo "== $ENVV $*"; env $ENVV $T /tmp/c1.mtx "$@" --repeat 2000 | grep -E "kernel:|wrong|min time"; }
for N in 8 16 32; do
 for cap in 8 16 32 64; do ENVV="GS_MERGE_CAP=$cap" run $N --kernel b200_merge; done
 ENVV="X=1" run $N --ops "merge_path_thread_operator,32;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1"
 ENVV="X=1" run $N --ops "sort_operator;merge_path_thread_operator,32;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1"
 ENVV="X=1" run $N --half --kernel b200_merge
 ENVV="GS_MERGE_THREAD=0" run $N --kernel b200_merge
done
