for sk in 0 1 2 4 8 16 32 3 19 23 31; do echo "== skip $sk"; GS_TILE_SKIP=$sk timeout 100 python scratch/dbg_tile.py 2>&1 | tail -1; done
