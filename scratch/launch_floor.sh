#!/bin/bash
# per-launch floor of the timing loop: a tiny matrix, every kernel family
cd /tmp && python -c "
import sys; sys.path.insert(0,'/root/repo')
import numpy as np
from generalsparse_b200 import synth
r,c=synth.synth_matrix('uniform',256,256,seed=1,density=0.05); synth.write_mtx('/tmp/tiny.mtx',256,256,r,c)"
T=/root/repo/generalsparse_b200/lib/token_test
for k in b200_row_smem b200_merge b200_nnz b200_sp24 b200_tcgen05 reference; do
  h=""; case $k in b200_sp24|b200_tcgen05) h="--half";; esac
  echo "== $k"; $T /tmp/tiny.mtx 8 $h --kernel $k --repeat 20000 | grep -E "kernel:|min time"
done
