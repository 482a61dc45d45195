#!/bin/bash
# tile-local remainder at N = 64 (two stages, two CTAs per SM): tests, then C3's unstructured half and C2 at N = 64 against the tile kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "sp24" > gpurun_out/t60.log 2>&1; echo "sp24 tests exit $?"; tail -3 gpurun_out/t60.log | cut -c1-300
for dens in 0.1 0.3; do
for k in b200_sp24 b200_tcgen05; do
timeout 300 python bench.py --density $dens --dense-n 64 --kernel $k --steps 1000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b60_${k}_d$dens.err | tail -1 > gpurun_out/b60_${k}_d$dens.json
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/b60_${k}_d$dens.json").read()); print("$k density $dens N=64 %.2f us frac %.3f parity %s | %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["parity"]["ok"], d["config"]["kernel"][-70:]))
except Exception as e: print("$k $dens failed", e)
PY
done
done
