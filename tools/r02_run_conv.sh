#!/bin/bash
# same-box A/B of the warp-convergent TMA issuer (libocppo_conv.so = tc_gemm.cu built with -DOCPPO_TMA_CONV, the rest identical)
mkdir -p gpurun_out
L=oc_cleanrl_b200/csrc
cp $L/libocppo.so /tmp/lib_base.so; cp $L/libocppo_conv.so /tmp/lib_conv.so
for v in base conv base conv; do
  cp /tmp/lib_$v.so $L/libocppo.so
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/conv_bench_$v.json 2> gpurun_out/conv_bench_$v.err
  python - <<PY
import json
d = json.load(open("gpurun_out/conv_bench_$v.json"))
print("$v ms_per_step %.2f" % d["ms_per_step"], {k: round(x, 3) for k, x in d["kernel_ms_per_step"].items() if k.startswith("gemm")}, d["clocks"]["sm_mhz"], "loss", d["last_stats"]["loss"])
PY
done
for v in base conv; do
  cp /tmp/lib_$v.so $L/libocppo.so
  python tools/gemm_target.py 2 > gpurun_out/conv_plain_$v.log 2>&1 && ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -k regex:tc_gemm -s 12 -c 12 --csv --log-file gpurun_out/conv_ncu_$v.csv python tools/gemm_target.py 2 > gpurun_out/conv_ncu_$v.log 2>&1; echo "ncu $v rc=$?"
  python - <<PY
import csv
rows = list(csv.reader(l for l in open("gpurun_out/conv_ncu_$v.csv") if l.startswith('"')))
h = rows[0]; mi = h.index("Metric Name"); vi = h.index("Metric Value")
cyc = [r[vi].replace(",", "") for r in rows[1:] if r[mi] == "sm__cycles_elapsed.max"]
ns = [r[vi].replace(",", "") for r in rows[1:] if r[mi] == "gpu__time_duration.sum"]
print("$v cycles:", " ".join(cyc), "sum %.3fM" % (sum(float(c) for c in cyc) / 1e6)); print("$v ns:", " ".join(ns))
PY
done
cp /tmp/lib_base.so $L/libocppo.so
