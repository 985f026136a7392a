#!/bin/bash
# bring-up build: what bounds the forward GEMMs' epilogue?  ncu cycles of the second minibatch's twelve GEMMs with
# OCPPO_TC_DBG = 0 (production), 2 (no bulk tensor stores), 4096 (no ReLU decision bits), 4098 (neither)
mkdir -p gpurun_out
for v in 0 8192 2; do
  OCPPO_TC_DBG=$v python tools/gemm_target.py 2 > gpurun_out/epj_plain_$v.log 2>&1 && OCPPO_TC_DBG=$v ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -k regex:tc_gemm -s 12 -c 12 --csv --log-file gpurun_out/epj_ncu_$v.csv python tools/gemm_target.py 2 > gpurun_out/epj_ncu_$v.log 2>&1; echo "ncu $v rc=$?"
  python - <<PY
import csv
rows = list(csv.reader(l for l in open("gpurun_out/epj_ncu_$v.csv") if l.startswith('"')))
h = rows[0]; mi = h.index("Metric Name"); vi = h.index("Metric Value")
cyc = [r[vi].replace(",", "") for r in rows[1:] if r[mi] == "sm__cycles_elapsed.max"]
print("dbg=$v cycles:", " ".join(cyc), "sum %.3fM" % (sum(float(c) for c in cyc) / 1e6))
PY
done
