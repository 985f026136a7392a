#!/bin/bash
for d in 0 16 32 48; do
OCPPO_TC_DBG=$d timeout 300 python bench.py --gemm tc-f16 --no-cpu-baseline --no-e2e --steps 2 --gae-sweep 4096 > gpurun_out/bench_f16_$d.json 2> gpurun_out/bench_f16_$d.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_f16_$d.json').read().strip().splitlines()[-1])
print('dbg=$d', d['ms_per_step'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['power_w_max'])
"
done
