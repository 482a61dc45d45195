#!/bin/bash
cd /root/repo
for N in 8 32; do
for e in "X=1" "GS_SP24_SKIP_REM=1" "ISO_PATTERN=nm24" "GS_SP24_FUSED_REM=0" "GS_SP24_FUSED_REM=1"; do
  env $e ISO_N=$N TAG="N=$N $e" python scratch/iso_sp24.py 2>&1 | grep "iso median"
done; done
