#!/bin/bash
# tile-local remainder of the 2:4 kernel: parity tests, then C2 at N = 8 / 16 / 32 against the previous modes
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "sp24" > gpurun_out/t43.log 2>&1; echo "sp24 tests exit $?"; tail -5 gpurun_out/t43.log
for n in 8 16 32; do
  timeout 300 python bench.py --kernel b200_sp24 --dense-n $n --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b43_n$n.err | tail -1 > gpurun_out/b43_n$n.json; echo "n$n exit $?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/b43_n$n.json").read()); print("N=$n", d["ms_per_step"]*1e3, "us", d["roofline"]["frac"], d["config"].get("kernel"), d.get("parity"), d.get("e2e"))
PY
done
GS_SP24_SKIP_REM=1 timeout 300 python bench.py --kernel b200_sp24 --dense-n 16 --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b43_n16_skip.err | tail -1 | cut -c1-200
GS_SP24_FUSED_REM=1 timeout 300 python bench.py --kernel b200_sp24 --dense-n 16 --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b43_n16_m1.err | tail -1 | cut -c1-200
timeout 600 ncu --metrics launch__occupancy_limit_shared_mem,launch__occupancy_limit_registers,launch__occupancy_limit_warps,launch__registers_per_thread,launch__shared_mem_per_block_dynamic,launch__shared_mem_per_block_static,launch__shared_mem_config_size,gpu__time_duration.sum --clock-control none -k regex:k_b200_sp24 -s 20 -c 2 --csv --log-file gpurun_out/occ43.csv python bench.py --kernel b200_sp24 --dense-n 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/occ43.log 2>&1; echo "ncu exit $?"
