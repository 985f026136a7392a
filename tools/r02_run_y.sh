#!/bin/bash
# bring-up build A/B: skinny4 load depth 4 / 8 (cuda_core class), with the thin-forward and loss-backward prefetches in
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_graphs.py -m gpu -x -q 2>&1 | tail -2
for v in 4 8 4 8; do
  OCPPO_SK4_DEPTH=$v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/y_bench_$v.json 2> gpurun_out/y_bench_$v.err
  python - <<PY
import json
d = json.load(open("gpurun_out/y_bench_$v.json"))
print("depth=$v ms_per_step %.2f" % d["ms_per_step"], {k: round(x, 3) for k, x in d["kernel_ms_per_step"].items()}, d["clocks"]["sm_mhz"])
PY
done
