#!/bin/bash
# timing experiments on the FP16x3 engine (numerics are wrong under these knobs): 16 = no conversion, 32 = main product only
mkdir -p gpurun_out
for d in ${DBGS:-0 16 32 48}; do
OCPPO_TC_DBG=$d timeout 300 python bench.py --gemm tc --no-cpu-baseline --no-e2e --gae-sweep 4096 > gpurun_out/bench_dbg$d.json 2> gpurun_out/bench_dbg$d.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_dbg$d.json').read().strip().splitlines()[-1])
print('dbg$d', d['ms_per_step'], {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()}, d['clocks'])
"
done
