#!/bin/bash
# after the tile-local remainder: whole GPU suite, smoke, the default bench line and N = 16 / 32, ncu traffic of the default command, launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t46.log 2>&1; echo "gpu tests exit $?"; tail -3 gpurun_out/t46.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/b46_default.json 2> gpurun_out/b46_default.err; echo "default exit $?"; cut -c1-400 gpurun_out/b46_default.json
for n in 16 32; do
  timeout 300 python bench.py --dense-n $n --no-cpu-baseline 2> gpurun_out/b46_n$n.err | tail -1 > gpurun_out/b46_n$n.json; echo "n$n exit $?"; cut -c1-330 gpurun_out/b46_n$n.json
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_b200_sp24 -s 20 -c 2 -f -o gpurun_out/sp24_tl_n8 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu46_n8.log 2>&1; echo "ncu n8 exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches46.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ncu46_list.log 2>&1; echo "ncu list exit $?"
