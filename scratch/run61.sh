#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t61.log 2>&1; echo "gpu tests exit $?"; tail -2 gpurun_out/t61.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 300 python bench.py --density 0.1 --dense-n 64 --steps 1000 --warmup 20 --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-420
timeout 400 python bench.py 2>/dev/null | tail -1 | cut -c1-200
