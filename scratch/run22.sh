export GS_TILE_CONSUMER_FENCE=1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_tile -s 4 -c 1 -o gpurun_out/prof_r1_tile_v4 -f python bench.py --n 32 --kernel b200_tcgen05 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tile_v4.log 2>&1; tail -1 gpurun_out/ncu_tile_v4.log
