#!/bin/bash
# first contact of the 2:4 kernel: probe, then its tests, then first timings
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 300 python scratch/probe_sp24.py > gpurun_out/probe_sp24.log 2>&1
echo "probe exit $?" >> gpurun_out/probe_sp24.log
tail -40 gpurun_out/probe_sp24.log
timeout 900 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "sp24" > gpurun_out/sp24_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/sp24_tests.log
tail -15 gpurun_out/sp24_tests.log
for n in 64 128 256; do
  timeout 300 python bench.py --pattern nm24 --n $n --steps 1000 --warmup 20 --no-cpu-baseline > gpurun_out/sp24_nm24_n$n.json 2> gpurun_out/sp24_nm24_n$n.err
  echo "nm24 n=$n exit $?"; cut -c1-400 gpurun_out/sp24_nm24_n$n.json
done
for n in 16 32; do
  timeout 300 python bench.py --kernel b200_sp24 --n $n --steps 1000 --warmup 20 --no-cpu-baseline > gpurun_out/sp24_c2_n$n.json 2> gpurun_out/sp24_c2_n$n.err
  echo "c2 n=$n exit $?"; cut -c1-400 gpurun_out/sp24_c2_n$n.json
done
