J='import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["config"]["kernel"], d["parity"])'
for n in 16 32; do echo "== N=$n"; timeout 300 python bench.py --n $n --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"; done
echo "== N=32 no pdl"; GS_NO_PDL=1 timeout 300 python bench.py --n 32 --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"
echo "== N=32 no bank order"; GS_TILE_NO_BANK_ORDER=1 timeout 300 python bench.py --n 32 --kernel b200_tcgen05 --steps 2000 --warmup 20 --no-cpu-baseline 2>&1 | tail -1 | python -c "$J"
timeout 600 python -m pytest tests/test_spmm_gpu.py -x -q -m gpu -k "tile" 2>&1 | tail -5
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_tile -s 4 -c 2 -o gpurun_out/prof_r1_tile_v1 -f python bench.py --n 32 --kernel b200_tcgen05 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tile_v1.log 2>&1; tail -3 gpurun_out/ncu_tile_v1.log
