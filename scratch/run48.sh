#!/bin/bash
# the driver's two-rank launch of the default bench line (now the 2:4 kernel with the tile-local remainder) and N=16
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541"
timeout 400 $TR bench.py --gpus 2 --steps 2000 --warmup 20 2> gpurun_out/b48_g2.err | tail -1 > gpurun_out/b48_g2.json; echo "g2 exit $?"; cut -c1-300 gpurun_out/b48_g2.json
timeout 400 $TR bench.py --gpus 2 --dense-n 32 --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b48_n32_g2.err | tail -1 > gpurun_out/b48_n32_g2.json; echo "n32 g2 exit $?"; cut -c1-300 gpurun_out/b48_n32_g2.json
