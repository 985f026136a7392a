#!/bin/bash
# BASELINE configs[2]: the learner step on SpaceInvaders- / Breakout-shaped observations (device-resident timing only)
mkdir -p gpurun_out
for g in ${GAMES:-spaceinvaders breakout}; do
timeout 600 python bench.py --game $g --no-cpu-baseline --no-e2e --gae-sweep 4096 > gpurun_out/bench_$g.json 2> gpurun_out/bench_$g.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_$g.json').read().strip().splitlines()[-1])
print('$g', round(d['value']), round(d['ms_per_step'],2), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()}, d['clocks']['sm_mhz'], 'inf_ms', [round(i['ms_per_call'],3) for i in d['inference']])
"
tail -2 gpurun_out/bench_$g.err
done
