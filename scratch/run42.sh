#!/bin/bash
# the driver's multi-rank launch of the default bench line and of the reference arm, two ranks
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 400 $TR bench.py --gpus 2 --steps 2000 --warmup 20 2> gpurun_out/bench_g2.err | tail -1 > gpurun_out/bench_g2.json; echo "g2 exit $?"; cut -c1-300 gpurun_out/bench_g2.json
timeout 400 $TR bench.py --impl reference --gpus 2 --steps 10 --warmup 3 2> gpurun_out/bench_ref_g2.err | tail -1 > gpurun_out/bench_ref_g2.json; echo "ref g2 exit $?"; cut -c1-200 gpurun_out/bench_ref_g2.json
timeout 400 $TR bench.py --gpus 2 --dense-n 16 --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/bench_n16_g2.err | tail -1 > gpurun_out/bench_n16_g2.json; echo "n16 g2 exit $?"; cut -c1-300 gpurun_out/bench_n16_g2.json
