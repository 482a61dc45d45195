#!/bin/bash
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
from generalsparse_b200 import synth
rows,cols,r,c=synth.c1_standin(); synth.write_mtx('/tmp/c1.mtx',rows,cols,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
cd /root/repo
ncu --set full --clock-control none --import-source on -k regex:k_b200_merge -s 5 -c 1 -o gpurun_out/r2_c1_chunk $T /tmp/c1.mtx 8 --repeat 20 --kernel b200_merge > gpurun_out/r2_c1_chunk_ncu.log 2>&1; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_b200_merge -s 5 -c 10 --csv --log-file gpurun_out/r2_c1_chunk_launches.csv $T /tmp/c1.mtx 8 --repeat 20 --kernel b200_merge > /dev/null 2>&1; echo rc=$?
