timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/bench_v4.json 2> gpurun_out/bench_v4.err; tail -c 3000 gpurun_out/bench_v4.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_stream -s 30 -c 1 -o gpurun_out/prof_r1_stream_v4 -f python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_v4.log 2>&1; tail -2 gpurun_out/ncu_v4.log
