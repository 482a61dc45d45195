#!/bin/bash
# ncu --set full of the tile-local 2:4 kernel at N = 32 and N = 16 on C2 (one launch each, source counters)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for n in 32 16; do
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_b200_sp24 -s 20 -c 1 -f -o gpurun_out/sp24_tl_n$n python bench.py --kernel b200_sp24 --dense-n $n --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu45_n$n.log 2>&1; echo "ncu n$n exit $?"
done
ls -la gpurun_out/*.ncu-rep
