timeout 600 python bench.py --workload c4 --rows 2000000 --steps 20 --warmup 3 2>&1 | tail -2
