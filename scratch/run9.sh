timeout 300 python scratch/probe_tile.py 2>&1 | tail -60
