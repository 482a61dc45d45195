#!/bin/bash
# compute-sanitizer memcheck over the tile-local remainder tests of the 2:4 kernel (small matrices)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 77 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "tile_local_remainder" > gpurun_out/san49.log 2>&1; echo "memcheck exit $?"
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds" gpurun_out/san49.log | head -10
