#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t57.log 2>&1; echo "gpu tests exit $?"; tail -2 gpurun_out/t57.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 1500 python scratch/sweep_c5.py > gpurun_out/c5_57.log 2>&1; echo "sweep exit $?"
grep "b200_sp24\|b200_row_smem\|b200_tcgen05\|^###" gpurun_out/c5_sweep.md | cut -c1-200
