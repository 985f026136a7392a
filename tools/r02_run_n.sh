#!/bin/bash
# final evidence pass of round 2 on one B200 (release build): whole GPU suite, smoke, the driver's bench line, ncu launch list + GEMM minibatch capture
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -x -q -m gpu ) > gpurun_out/n_pytest_all.txt 2>&1; echo "all (-x) rc=$?"; grep -E "passed|failed" gpurun_out/n_pytest_all.txt
( time python __graft_entry__.py --smoke ) > gpurun_out/n_smoke.txt 2>&1; echo "smoke rc=$?"; grep smoke gpurun_out/n_smoke.txt
( time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/n_bench.json ) 2> gpurun_out/n_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/n_bench.json"))
print("ms_per_step %.2f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), "frac %.4f" % d["roofline"]["frac"], {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
PY
CMD="python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep 4096"
$CMD > gpurun_out/n_plain.json 2> gpurun_out/n_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/n_launches.csv $CMD > gpurun_out/n_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python tools/gemm_target.py 2 > gpurun_out/n_gemm_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 12 -c 12 -o gpurun_out/n_gemm_prof python tools/gemm_target.py 2 > gpurun_out/n_gemm_ncu.log 2>&1; echo "ncu gemm rc=$?"
