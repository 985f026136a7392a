#!/bin/bash
# bring-up build A/B: register-resident loss-head backward on / off (heads class of the profiled pass)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_graphs.py -m gpu -x -q 2>&1 | tail -2
for v in 1 0 1 0; do
  OCPPO_LOSS_BWD_REG=$v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/w_bench_$v.json 2> gpurun_out/w_bench_$v.err
  python - <<PY
import json
d = json.load(open("gpurun_out/w_bench_$v.json"))
print("reg=$v ms_per_step %.2f" % d["ms_per_step"], {k: round(x, 3) for k, x in d["kernel_ms_per_step"].items()}, d["clocks"]["sm_mhz"])
PY
done
