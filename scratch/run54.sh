#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t54.log 2>&1; echo "gpu tests exit $?"; tail -2 gpurun_out/t54.log
