G=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --workload c4 --gpus $G --steps 20 --warmup 3 2> gpurun_out/c4_g$G.err | tail -1 > gpurun_out/c4_g$G.json; cut -c1-400 gpurun_out/c4_g$G.json
timeout 600 $TR bench.py --gpus $G --steps 2000 --warmup 20 2> gpurun_out/c2_n8_g$G.err | tail -1 > gpurun_out/c2_n8_g$G.json; cut -c1-300 gpurun_out/c2_n8_g$G.json
timeout 600 $TR bench.py --gpus $G --n 32 --steps 2000 --warmup 20 2> gpurun_out/c2_n32_g$G.err | tail -1 > gpurun_out/c2_n32_g$G.json; cut -c1-300 gpurun_out/c2_n32_g$G.json
