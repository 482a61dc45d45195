GS_ROW_SCHED=stream timeout 900 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "b200 or row" 2>&1 | tail -3
J='import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["roofline"]["frac"], d["config"]["kernel"], d["parity"]["ok"])'
run() { echo "== $*"; env "$@" timeout 300 python bench.py --steps 3000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"; }
run X=1
run GS_STREAM_CTAS=3
run GS_STREAM_CTAS=3 GS_STREAM_WARPS=9
run GS_STREAM_CTAS=1
run GS_STREAM_PF=2
run GS_BENCH_ROWS=44032
