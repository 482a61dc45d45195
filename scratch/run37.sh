#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for n in 8 16 32; do
  timeout 300 python bench.py --kernel b200_sp24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v6_c2_n$n.json 2> gpurun_out/sp24v6_c2_n$n.err
  echo "c2 n=$n exit $?"; cut -c1-220 gpurun_out/sp24v6_c2_n$n.json
done
timeout 600 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "sp24" > gpurun_out/sp24_tests6.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/sp24_tests6.log
timeout 900 python scratch/sweep_c5.py > gpurun_out/c5_sweep.log 2>&1; echo "sweep exit $?"; tail -30 gpurun_out/c5_sweep.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke2.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke2.log | cut -c1-300
