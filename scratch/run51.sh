#!/bin/bash
# gs_spmm_host_async with two alternating launch streams: tests of the host-buffer paths, then the e2e arm of the bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "host or async or sp24" > gpurun_out/t51.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/t51.log
show() {
  python - <<PY
import json
d=json.loads(open("gpurun_out/b51_$1.json").read()); e=d["e2e"]; print("$1 %.2f us | e2e async %.2f us, host enqueue %.2f us, blocking %.2f us | %s" % (d["ms_per_step"]*1e3, e["ms_per_step"]*1e3, e["host_enqueue_ms_per_step"]*1e3, e["blocking_ms_per_step"]*1e3, e["api"][-30:]))
PY
}
timeout 300 python bench.py --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b51_sp24.err | tail -1 > gpurun_out/b51_sp24.json; show sp24
GS_ASYNC_ONE_STREAM=1 timeout 300 python bench.py --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b51_sp24_one.err | tail -1 > gpurun_out/b51_sp24_one.json; show sp24_one
timeout 300 python bench.py --kernel b200_row_smem --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b51_row.err | tail -1 > gpurun_out/b51_row.json; show row
timeout 300 python bench.py --dense-n 32 --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b51_n32.err | tail -1 > gpurun_out/b51_n32.json; show n32
timeout 300 python bench.py --dense-n 64 --density 0.1 --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b51_c3.err | tail -1 > gpurun_out/b51_c3.json; show c3
