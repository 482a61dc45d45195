GS_TILE_DEBUG=1 timeout 300 python scratch/dbg_tile.py 2>&1 | tail -17 | cut -c1-22,88-200
