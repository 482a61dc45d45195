#!/bin/bash
# chunk kernel on the C1 stand-in: staging x tickets x cap x N x dtype on ONE box
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
from generalsparse_b200 import synth
rows,cols,r,c=synth.c1_standin(); synth.write_mtx('/tmp/c1.mtx',rows,cols,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
for h in "" "--half"; do for N in 8 16 32; do for cap in 32 64; do
  line="N=$N $h cap=$cap:"
  for st in 0 1; do for tk in 0 1; do
    t=$(GS_MERGE_CAP=$cap GS_MERGE_STAGE=$st GS_MERGE_TICKET=$tk $T /tmp/c1.mtx $N $h --kernel b200_merge --repeat 3000 | grep -E "min time|wrong" | sed 's/.*min time: //' | tr '\n' ' ')
    line="$line  stage$st/ticket$tk=$t"
  done; done
  echo "$line"
done; done; done
echo "row kernel:"; for N in 8 16 32; do $T /tmp/c1.mtx $N --kernel b200_row_smem --repeat 3000 | grep -E "kernel:|min time" | paste - - | cut -c1-150; done
echo "auto:"; for N in 8 16 32; do $T /tmp/c1.mtx $N --repeat 3000 | grep -E "kernel:|min time" | paste - - | cut -c1-200; done
