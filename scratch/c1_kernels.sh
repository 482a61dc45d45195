#!/bin/bash
# C1 stand-in on every kernel that takes it: GFLOP/s and time per launch from the token_test driver
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
from generalsparse_b200 import synth
rows,cols,r,c=synth.c1_standin(); synth.write_mtx('/tmp/c1.mtx',rows,cols,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
run() { echo "== $*"; $T /tmp/c1.mtx "$@" --repeat 2000 | grep -E "kernel:|correct|wrong|min time|c1.mtx [0-9]"; }
for N in 8 32; do
run $N --kernel reference
run $N --kernel auto
run $N --sparse-coarsen 8 --kernel b200_row_smem
run $N --kernel b200_merge
for W in 16 32 64 128 256; do
run $N --ops "merge_path_thread_operator,$W;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1"
done
run $N --ops "sort_operator;merge_path_thread_operator,64;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1"
run $N --ops "merge_path_warp_operator,1024;warp_total_reduce_operator,1;grid_block_operator,64,32,4,1"
run $N --ops "balanced_interval_row_direction_thread_blocking_operator,64,0,0;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1"
run $N --family warp_bitmap --kernel b200_merge
run $N --family warp_segment --kernel reference
done
run 32 --family warp_segment --kernel b200_nnz
run 8 --half --kernel b200_merge
run 8 --half --kernel auto
