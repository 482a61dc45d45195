for sk in 0 32 33 2 4 34 38 39; do echo "== skip $sk"; GS_TILE_SKIP=$sk timeout 100 python scratch/dbg_tile.py 2>&1 | tail -1; done
