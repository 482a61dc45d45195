GS_TILE_DEBUG=1 timeout 300 python scratch/dbg_tile.py 2>&1 | grep -A40 "per-tile trace" | head -42
