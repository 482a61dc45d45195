#!/bin/bash
# C3's unstructured half (90 % sparse) at N = 64 / 128 / 256: 2:4 image + gather pass against the tile kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for n in 64 128 256; do
for k in b200_sp24 b200_tcgen05; do
timeout 300 python bench.py --density 0.1 --dense-n $n --kernel $k --steps 1000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b59_${k}_n$n.err | tail -1 > gpurun_out/b59_${k}_n$n.json
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/b59_${k}_n$n.json").read()); print("$k N=$n %.2f us frac %.3f parity %s | %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["parity"]["ok"], d["config"]["kernel"][-60:]))
except Exception as e: print("$k N=$n failed", e)
PY
done
done
