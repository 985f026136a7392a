#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "gae" ) > gpurun_out/j_pytest_gae.txt 2>&1; echo "gae tests rc=$?"; tail -2 gpurun_out/j_pytest_gae.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/j_bench.json 2> gpurun_out/j_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/j_bench.json"))
print("ms_per_step %.2f" % d["ms_per_step"], {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"]["sm_mhz"], "loss", d["last_stats"]["loss"])
PY
( timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "golden or option_switches or chunked" ) > gpurun_out/j_pytest_loss.txt 2>&1; echo "loss tests rc=$?"; tail -2 gpurun_out/j_pytest_loss.txt
