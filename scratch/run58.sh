#!/bin/bash
# the driver's 8-rank launch of the default bench line (2:4 kernel, tile-local remainder) and of the reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551"
timeout 500 $TR bench.py --gpus 8 --steps 2000 --warmup 20 2> gpurun_out/b58_g8.err | tail -1 > gpurun_out/b58_g8.json; echo "g8 exit $?"; cut -c1-260 gpurun_out/b58_g8.json
timeout 300 $TR bench.py --impl reference --gpus 8 --steps 5 --warmup 3 2> gpurun_out/b58_ref_g8.err | tail -1 > gpurun_out/b58_ref_g8.json; echo "ref g8 exit $?"; cut -c1-200 gpurun_out/b58_ref_g8.json
