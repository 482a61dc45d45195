#!/bin/bash
cd /root/repo
for N in 8 32; do
for e in "X=1" "GS_SP24_PREFETCH=0" "GS_SP24_PREFETCH=8" "GS_SP24_PREFETCH=16" "GS_SP24_PREFETCH=32" "GS_SP24_TL_STAGES=3" "GS_TILE_CTAS=296" "GS_TILE_CTAS=296 GS_SP24_PREFETCH=16" "GS_NO_PDL=1" "GS_SP24_SKIP_EMPTY=0"; do
  env $e ISO_N=$N TAG="N=$N $e" python scratch/iso_sp24.py 2>&1 | grep "iso median"
done; done
