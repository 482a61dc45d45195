#!/bin/bash
# C4 kernel probes on one B200: parity tests of the nnz kernel, timing variants, then one ncu capture
python -m pytest tests -m gpu -x -q -k "nnz" 2>&1 | tail -25 > gpurun_out/nnz_tests.log
tail -3 gpurun_out/nnz_tests.log
for v in "X=0" "GS_NNZ_OCC=5" "GS_NNZ_OCC=3" "GS_NNZ_GROUP=16" "GS_NNZ_GROUP=4" "GS_NNZ_RUN=4" "GS_NNZ_RUN=16" "GS_NNZ_RUN=31"; do
  env $v python bench.py --workload c4 --steps 20 --warmup 3 2>&1 | tail -1 > "gpurun_out/c4_probe_$v.json"
  echo "$v $(python -c "import json,sys; d=json.loads(open('gpurun_out/c4_probe_$v.json').read()); print(d['ms_per_step'], d['parity'])")"
done
ncu --set full --clock-control none --import-source on -k regex:k_b200_nnz --launch-skip 3 -c 1 -o gpurun_out/c4_nnz python bench.py --workload c4 --steps 3 --warmup 3 > gpurun_out/c4_ncu.log 2>&1
ls -la gpurun_out/
