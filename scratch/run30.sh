set -x
S=$(date +%s)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "pytest wall $(( $(date +%s) - S )) s"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/r1_bench_n8.json 2> gpurun_out/r1_bench_n8.err; tail -1 gpurun_out/r1_bench_n8.json | cut -c1-1500
for N in 16 32; do
  timeout 600 python bench.py --n $N --no-cpu-baseline > gpurun_out/r1_bench_n$N.json 2> gpurun_out/r1_bench_n$N.err; tail -1 gpurun_out/r1_bench_n$N.json | cut -c1-1200
done
for N in 64 128 256; do
  timeout 600 python bench.py --density 0.10 --n $N --no-cpu-baseline > gpurun_out/r1_bench_c3_n$N.json 2> gpurun_out/r1_bench_c3_n$N.err; tail -1 gpurun_out/r1_bench_c3_n$N.json | cut -c1-1200; tail -2 gpurun_out/r1_bench_c3_n$N.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_tile_n32.csv python bench.py --n 32 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ll.log 2>&1; tail -2 gpurun_out/ncu_ll.log | cut -c1-300
for N in 16 32; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_b200_tile --launch-skip 12 -c 2 -o gpurun_out/prof_r1_tile_final_n$N -f python bench.py --n $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_n$N.log 2>&1; tail -2 gpurun_out/ncu_full_n$N.log | cut -c1-300
done
echo "total wall $(( $(date +%s) - S )) s"
