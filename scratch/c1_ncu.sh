#!/bin/bash
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
from generalsparse_b200 import synth
rows,cols,r,c=synth.c1_standin(); synth.write_mtx('/tmp/c1.mtx',rows,cols,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
cd /root/repo
ncu --set full --clock-control none --import-source on -k regex:k_b200_merge -s 5 -c 1 -o gpurun_out/r2_c1_merge $T /tmp/c1.mtx 8 --repeat 20 --ops "merge_path_thread_operator,32;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1" > gpurun_out/r2_c1_merge_ncu.log 2>&1; echo rc=$?
GS_ROW_GROUP=0 ncu --set full --clock-control none --import-source on -k regex:k_b200_row -s 5 -c 1 -o gpurun_out/r2_c1_row $T /tmp/c1.mtx 8 --repeat 20 --kernel b200_row_smem > gpurun_out/r2_c1_row_ncu.log 2>&1; echo rc=$?
