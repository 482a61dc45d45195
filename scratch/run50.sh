#!/bin/bash
# e2e arm: host enqueue time against the whole round trip, 2:4 kernel and gather kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for k in b200_sp24 b200_row_smem; do
  timeout 300 python bench.py --kernel $k --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b50_$k.err | tail -1 > gpurun_out/b50_$k.json
  python - <<PY
import json
d=json.loads(open("gpurun_out/b50_$k.json").read()); e=d["e2e"]; print("$k %.2f us | e2e async %.2f us, host enqueue %.2f us, blocking %.2f us" % (d["ms_per_step"]*1e3, e["ms_per_step"]*1e3, e["host_enqueue_ms_per_step"]*1e3, e["blocking_ms_per_step"]*1e3))
PY
done
