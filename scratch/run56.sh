#!/bin/bash
# 2:4 image without its empty tiles: sp24 tests, default bench (no empty tiles: must not move), C5's banded matrix
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "sp24" > gpurun_out/t56.log 2>&1; echo "sp24 tests exit $?"; tail -4 gpurun_out/t56.log | cut -c1-300
for n in 8 32; do
timeout 300 python bench.py --dense-n $n --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b56_n$n.err | tail -1 > gpurun_out/b56_n$n.json
python - <<PY
import json
d=json.loads(open("gpurun_out/b56_n$n.json").read()); print("N=$n %.2f us frac %.3f parity %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["parity"]["ok"]))
PY
done
timeout 300 python bench.py --pattern nm24 --dense-n 16 --steps 2000 --warmup 20 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('nm24 N=16 %.2f us frac %.3f' % (d['ms_per_step']*1e3, d['roofline']['frac']))"
