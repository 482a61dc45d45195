#!/bin/bash
# C5 sweep again with the 2:4 kernel as one more row
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python scratch/sweep_c5.py > gpurun_out/c5_55.log 2>&1; echo "sweep exit $?"; tail -3 gpurun_out/c5_55.log | cut -c1-200
