#!/bin/bash
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -x -q -m gpu ) > gpurun_out/o_pytest_all.txt 2>&1; echo "all (-x) rc=$?"; grep -E "passed|failed" gpurun_out/o_pytest_all.txt
( time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/o_bench.json ) 2> gpurun_out/o_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/o_bench.json"))
print("ms_per_step %.2f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), "frac %.4f" % d["roofline"]["frac"], {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
PY
