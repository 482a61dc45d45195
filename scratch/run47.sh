#!/bin/bash
# one launch alone (copies on both sides, the e2e arm) against a stream of launches: 148 CTAs (one per SM per launch, two launches overlap) or 296 (two per SM)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
run() {
  local label=$1; shift
  env "$@" timeout 300 python bench.py --dense-n $N --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b47_$label.err | tail -1 > gpurun_out/b47_$label.json
  python - <<PY
import json
d=json.loads(open("gpurun_out/b47_$label.json").read()); e=d["e2e"]; print("$label N=$N %.2f us frac %.3f parity %s e2e async %.2f us blocking %.2f us" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["parity"]["ok"], e["ms_per_step"]*1e3, e["blocking_ms_per_step"]*1e3))
PY
}
N=8 run n8_148 GS_TILE_CTAS=148
N=8 run n8_296 GS_TILE_CTAS=296
N=8 run n8_222 GS_TILE_CTAS=222
N=16 run n16_296 GS_TILE_CTAS=296
N=8 timeout 300 python bench.py --kernel b200_row_smem --dense-n 8 --steps 2000 --warmup 20 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['e2e']; print('gather N=8 %.2f us e2e async %.2f blocking %.2f' % (d['ms_per_step']*1e3, e['ms_per_step']*1e3, e['blocking_ms_per_step']*1e3))"
