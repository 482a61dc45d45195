#!/bin/bash
# final state of the round: whole GPU suite, smoke, default bench line, reference arm, N=16/32 lines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t53.log 2>&1; echo "gpu tests exit $?"; tail -2 gpurun_out/t53.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/b53_default.json 2> gpurun_out/b53_default.err; echo "default exit $?"
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/b53_reference.json 2> gpurun_out/b53_reference.err; echo "reference exit $?"; cut -c1-200 gpurun_out/b53_reference.json
for n in 16 32; do
  timeout 300 python bench.py --dense-n $n --no-cpu-baseline 2> gpurun_out/b53_n$n.err | tail -1 > gpurun_out/b53_n$n.json; echo "n$n exit $?"
done
python - <<'PY'
import json
for f in ("default","n16","n32"):
    d=json.loads(open("gpurun_out/b53_%s.json"%f).read().strip().splitlines()[-1]); e=d["e2e"]
    print(f, "%.2f us frac %.3f | e2e %.2f us (%.0f GFLOP/s) enqueue %.2f | cpu %s | clocks %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], e["ms_per_step"]*1e3, e["value"], e["host_enqueue_ms_per_step"]*1e3, (d.get("cpu_baseline") or {}).get("value"), d["clocks"]))
PY
