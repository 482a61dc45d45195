#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "sp24" > gpurun_out/sp24_tests7.log 2>&1; echo "tests exit $?"; tail -12 gpurun_out/sp24_tests7.log | cut -c1-200
for n in 8 16; do
  timeout 200 python bench.py --kernel b200_sp24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v7_c2_n$n.json 2> gpurun_out/sp24v7_c2_n$n.err
  echo "c2 n=$n exit $?"; cut -c1-220 gpurun_out/sp24v7_c2_n$n.json; tail -2 gpurun_out/sp24v7_c2_n$n.err | cut -c1-200
done
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke3.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke3.log | cut -c1-200
