#!/bin/bash
# C4 strong scaling of the nnz kernel on one 8-GPU box: 8, 4, 2 ranks (1 rank: profiles/r1_c4_nnz_g1.json)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for G in 8 4 2; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2951$G"
  timeout 400 $TR bench.py --workload c4 --gpus $G --steps 20 --warmup 3 2> gpurun_out/c4_nnz_g$G.err | tail -1 > gpurun_out/c4_nnz_g$G.json
  echo "c4 g=$G exit $?"; cut -c1-300 gpurun_out/c4_nnz_g$G.json
done
