#!/bin/bash
# where the tile-local remainder's time goes: no remainder math, stage count, prefetch distance
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
run() { # label, env..., -- n
  local label=$1; shift
  env "$@" timeout 300 python bench.py --kernel b200_sp24 --dense-n $N --steps 2000 --warmup 20 --no-cpu-baseline 2> gpurun_out/b44_$label.err | tail -1 > gpurun_out/b44_$label.json
  python - <<PY
import json
d=json.loads(open("gpurun_out/b44_$label.json").read()); print("$label N=$N %.2f us frac %.3f parity %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["parity"]["ok"]))
PY
}
N=32 run n32_skip GS_SP24_SKIP_REM=1
N=8 run n8_skip GS_SP24_SKIP_REM=1
N=32 run n32_pf0 GS_SP24_PREFETCH=0
N=32 run n32_pf8 GS_SP24_PREFETCH=8
N=16 run n16_s3 GS_SP24_TL_STAGES=3
N=16 run n16_pf8 GS_SP24_PREFETCH=8
N=16 run n16_pf0 GS_SP24_PREFETCH=0
N=8 run n8_s3 GS_SP24_TL_STAGES=3
