timeout 600 python -m pytest tests/test_spmm_gpu.py -q -m gpu -k "tile" 2>&1 | grep -E "^E  .*Error|passed|failed" | head -20
J='import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["config"]["kernel"], d["parity"])'
for n in 16 32; do echo "== N=$n"; timeout 300 python bench.py --n $n --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"; done
echo "== N=32 consumer fence"; GS_TILE_CONSUMER_FENCE=1 timeout 300 python bench.py --n 32 --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"
echo "== N=32 bulk zero"; GS_TILE_BULK_ZERO=1 timeout 300 python bench.py --n 32 --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"
GS_TILE_DEBUG=1 timeout 300 python scratch/dbg_tile.py 2>&1 | grep -A40 "per-tile trace" | awk 'NR%3==1' | head -16
