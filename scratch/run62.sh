#!/bin/bash
# launch list of the default command on the final code of the round
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b62.json 2> gpurun_out/b62.err; echo "bench exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches62.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu62.log 2>&1; echo "ncu list exit $?"
cut -d, -f5 gpurun_out/launches62.csv | sort | uniq -c | sort -rn | head -4
