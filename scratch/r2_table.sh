#!/bin/bash
# the rows of DESIGN.md section 5 / BASELINE.md section 3: one bench line per configuration (library auto selection), then the ncu
# tensor-pipe captures of the 2:4 kernel at N = 64 / 128 / 256
out=gpurun_out/r2_table.jsonl; : > $out
run() { python bench.py --no-c4 --no-cpu-baseline --no-yardstick --steps 1000 "$@" 2>/dev/null | grep '^{' >> $out; }
run --dense-n 8; run --dense-n 16; run --dense-n 32
run --dense-n 8 --kernel b200_row_smem; run --dense-n 16 --kernel b200_tcgen05; run --dense-n 32 --kernel b200_tcgen05
for n in 64 128 256; do run --dense-n $n --density 0.10; run --dense-n $n --pattern nm24; done
run --dense-n 16 --pattern nm24; run --dense-n 32 --pattern nm24
python - <<'PY'
import json
for l in open("gpurun_out/r2_table.jsonl"):
    d=json.loads(l); r=d["roofline"]
    print("%-52s %-30s us %7.2f frac %.3f fmt/alg %.2f dep %.2f iso %.2f e2e %.2f parity %s" % (d["config"]["workload"][28:80], d["config"]["kernel"][:30], d["ms_per_step"]*1e3, r["frac"], r["format_over_algorithmic"], r["dependent_chain_ms_per_step"]*1e3, r["isolated_launch_ms"]*1e3, d["e2e"]["ms_per_step"]*1e3, d["parity"]["ok"]))
PY
if [ -n "$R2_TABLE_NCU" ]; then for n in 64 128 256; do
  ncu --set full --clock-control none --import-source on -k regex:k_b200_sp24 -s 20 -c 1 -o gpurun_out/r2_sp24_nm24_n$n python bench.py --no-c4 --no-cpu-baseline --no-yardstick --steps 20 --warmup 5 --dense-n $n --pattern nm24 > gpurun_out/r2_sp24_ncu_n$n.log 2>&1; echo "ncu n$n rc=$?"
done; fi
