#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for n in 16 64 128 256; do
  for pf in 0 4; do
    GS_SP24_PREFETCH=$pf timeout 300 python bench.py --pattern nm24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v4_nm24_n${n}_pf$pf.json 2> gpurun_out/sp24v4_nm24_n${n}_pf$pf.err
    echo "nm24 n=$n pf=$pf exit $?"; cut -c1-220 gpurun_out/sp24v4_nm24_n${n}_pf$pf.json
  done
done
for n in 64 128; do
  GS_SP24_PREFETCH=0 GS_SP24_PANELS=2 timeout 300 python bench.py --pattern nm24 --n $n --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v4_nm24_n${n}_p2_pf0.json 2> gpurun_out/sp24v4_nm24_n${n}_p2_pf0.err
  echo "nm24 n=$n p2 pf0 exit $?"; cut -c1-220 gpurun_out/sp24v4_nm24_n${n}_p2_pf0.json
done
GS_SP24_PREFETCH=0 timeout 300 python bench.py --kernel b200_sp24 --n 8 --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/sp24v4_c2_n8_pf0.json 2> gpurun_out/sp24v4_c2_n8_pf0.err
echo "c2 n=8 pf0 exit $?"; cut -c1-220 gpurun_out/sp24v4_c2_n8_pf0.json
# the whole GPU suite, the smoke entry and the default bench line
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests_all.log 2>&1; echo "gpu tests exit $?"; tail -3 gpurun_out/gpu_tests_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_default.json
