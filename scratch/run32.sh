#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "sp24" > gpurun_out/sp24_tests2.log 2>&1
echo "tests exit $?" >> gpurun_out/sp24_tests2.log
tail -5 gpurun_out/sp24_tests2.log
for n in 16 32; do
  timeout 300 python bench.py --kernel b200_sp24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v2_c2_n$n.json 2> gpurun_out/sp24v2_c2_n$n.err
  echo "c2 n=$n exit $?"; cut -c1-300 gpurun_out/sp24v2_c2_n$n.json
  GS_SP24_SKIP_REM=1 timeout 300 python bench.py --kernel b200_sp24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v2_c2_n${n}_norem.json 2> gpurun_out/sp24v2_c2_n${n}_norem.err
  echo "c2 n=$n norem exit $?"; cut -c1-300 gpurun_out/sp24v2_c2_n${n}_norem.json
done
for n in 16 32 64 128 256; do
  for pp in 1 2; do
    GS_SP24_PANELS=$pp timeout 300 python bench.py --pattern nm24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v2_nm24_n${n}_p$pp.json 2> gpurun_out/sp24v2_nm24_n${n}_p$pp.err
    echo "nm24 n=$n panels=$pp exit $?"; cut -c1-300 gpurun_out/sp24v2_nm24_n${n}_p$pp.json
  done
done
# C4 on the nnz kernel, one GPU
timeout 600 python bench.py --workload c4 --steps 20 --warmup 3 > gpurun_out/c4_nnz_g1.json 2> gpurun_out/c4_nnz_g1.err
echo "c4 exit $?"; cut -c1-600 gpurun_out/c4_nnz_g1.json
# ncu: launch list and one full capture of the 2:4 kernel (N=64, 2:4 weight)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_sp24_n64.csv python bench.py --pattern nm24 --n 64 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_sp24 -s 20 -c 2 -o gpurun_out/sp24_n64_full python bench.py --pattern nm24 --n 64 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
