#!/bin/bash
# gs_spmm_host_async: 1 / 2 / 3 / 4 launch streams, 2:4 kernel and gather kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
show() {
  python - <<PY
import json
d=json.loads(open("gpurun_out/b52_$1.json").read()); e=d["e2e"]; print("$1 %.2f us | e2e async %.2f us, host enqueue %.2f us | equal %s" % (d["ms_per_step"]*1e3, e["ms_per_step"]*1e3, e["host_enqueue_ms_per_step"]*1e3, e["api"][-5:]))
PY
}
for n in 2 3 4; do
GS_ASYNC_STREAMS=$n timeout 300 python bench.py --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b52_sp24_s$n.err | tail -1 > gpurun_out/b52_sp24_s$n.json; show sp24_s$n
done
GS_ASYNC_STREAMS=3 timeout 300 python bench.py --kernel b200_row_smem --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b52_row_s3.err | tail -1 > gpurun_out/b52_row_s3.json; show row_s3
GS_ASYNC_STREAMS=3 timeout 300 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "host or async" 2>&1 | tail -1
