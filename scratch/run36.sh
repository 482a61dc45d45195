#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "sp24" > gpurun_out/sp24_tests5.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/sp24_tests5.log
for n in 128 256 64; do
  timeout 300 python bench.py --pattern nm24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v5_nm24_n${n}.json 2> gpurun_out/sp24v5_nm24_n${n}.err
  echo "nm24 n=$n exit $?"; cut -c1-220 gpurun_out/sp24v5_nm24_n${n}.json
done
