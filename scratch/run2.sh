GS_ROW_SCHED=stream timeout 900 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "b200 or row" 2>&1 | tail -3
J='import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["roofline"]["frac"], d["config"]["kernel"], d["parity"]["ok"])'
run() { echo "== $*"; env "$@" timeout 300 python bench.py --steps 3000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"; }
run X=1
run GS_STREAM_CTAS=1
run GS_ROW_NO_RAINBOW=1
run GS_STREAM_WARPS=14
run GS_STREAM_WARPS=12
run GS_STREAM_CTAS=1 GS_STREAM_WARPS=28
run GS_STREAM_CTAS=1 GS_STREAM_WARPS=24
run GS_BENCH_ROWS=44032
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_stream -s 30 -c 1 -o gpurun_out/prof_r1_stream_v3 -f python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_v3.log 2>&1; tail -2 gpurun_out/ncu_v3.log
