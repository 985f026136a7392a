#!/bin/bash
# iteration loop: GPU tests, then the bench (device-resident only) for each engine in $ENGINES under each OCPPO_TC_DBG in $DBGS
mkdir -p gpurun_out
[ -z "$NOTEST" ] && timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 --timeout-method thread 2>&1 | tail -6
for r in 1 2; do
for g in ${ENGINES:-tc}; do
for d in ${DBGS:-0}; do
OCPPO_TC_DBG=$d timeout 300 python bench.py --gemm $g --no-cpu-baseline --no-e2e --gae-sweep 4096 > gpurun_out/bench_${g}_d$d.json 2> gpurun_out/bench_${g}_d$d.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_${g}_d$d.json').read().strip().splitlines()[-1])
print('$g dbg$d', round(d['ms_per_step'],2), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()}, d['clocks']['sm_mhz'], d['clocks']['power_w_max'], d['last_stats']['loss'], d['last_stats']['grad_norm'], 'inf_ms', [round(i['ms_per_call'],3) for i in d['inference']])
"
tail -2 gpurun_out/bench_${g}_d$d.err
done
done
done
