timeout 600 python -m pytest tests/test_spmm_gpu.py -q -m gpu -k "tile" 2>&1 | grep -E "^E  .*Error|passed|failed" | head -20
J='import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["config"]["kernel"], d["parity"])'
for n in 16 32; do echo "== N=$n"; timeout 300 python bench.py --n $n --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"; done
GS_TILE_DEBUG=1 timeout 300 python scratch/dbg_tile.py 2>&1 | tail -17 | cut -c1-22,88-200
