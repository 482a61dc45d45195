#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "sp24" > gpurun_out/sp24_tests3.log 2>&1
echo "tests exit $?" >> gpurun_out/sp24_tests3.log
tail -5 gpurun_out/sp24_tests3.log
for n in 8 16 32; do
  timeout 300 python bench.py --kernel b200_sp24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v3_c2_n$n.json 2> gpurun_out/sp24v3_c2_n$n.err
  echo "c2 n=$n exit $?"; cut -c1-300 gpurun_out/sp24v3_c2_n$n.json
done
GS_SP24_SKIP_REM=1 timeout 300 python bench.py --kernel b200_sp24 --n 8 --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v3_c2_n8_norem.json 2> gpurun_out/sp24v3_c2_n8_norem.err
echo "c2 n=8 norem exit $?"; cut -c1-300 gpurun_out/sp24v3_c2_n8_norem.json
for n in 16 64 128 256; do
  timeout 300 python bench.py --pattern nm24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v3_nm24_n${n}.json 2> gpurun_out/sp24v3_nm24_n${n}.err
  echo "nm24 n=$n exit $?"; cut -c1-300 gpurun_out/sp24v3_nm24_n${n}.json
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_sp24_n64.csv python bench.py --pattern nm24 --n 64 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_sp24 -s 20 -c 2 -o gpurun_out/sp24_n64_full python bench.py --pattern nm24 --n 64 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
