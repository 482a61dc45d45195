J='import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["config"]["kernel"], d["parity"])'
for n in 16 32; do echo "== N=$n"; timeout 300 python bench.py --n $n --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"; done
echo "== N=16 row sched"; GS_ROW_SCHED=row timeout 300 python bench.py --n 16 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"
echo "== N=16 ctas1"; GS_STREAM_CTAS=1 timeout 300 python bench.py --n 16 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"
