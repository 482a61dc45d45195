#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests_all2.log 2>&1; echo "gpu tests exit $?"; tail -3 gpurun_out/gpu_tests_all2.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke4.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke4.log | cut -c1-400
timeout 600 python bench.py > gpurun_out/bench_default2.json 2> gpurun_out/bench_default2.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_default2.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench ref exit $?"; cut -c1-300 gpurun_out/bench_reference.json
timeout 300 python bench.py --n 16 --no-cpu-baseline > gpurun_out/bench_n16_auto.json 2> gpurun_out/bench_n16_auto.err; echo "n16 exit $?"; cut -c1-200 gpurun_out/bench_n16_auto.json
timeout 300 python bench.py --n 8 --kernel b200_sp24 --no-cpu-baseline > gpurun_out/bench_n8_sp24.json 2> gpurun_out/bench_n8_sp24.err; echo "n8 sp24 exit $?"; cut -c1-200 gpurun_out/bench_n8_sp24.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_sp24_c2_n16.csv python bench.py --n 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list2.log 2>&1; echo "ncu list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_sp24 -s 20 -c 2 -o gpurun_out/sp24_c2_n16_full python bench.py --n 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full2.log 2>&1; echo "ncu full exit $?"
