#!/bin/bash
# engine-5 iteration loop: GEMM tests of engine 5, the engine-5 parity tests, then the bench with both engines
timeout 300 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_parity.py -q -x -k "5] or f16" --timeout 150 --timeout-method thread 2>&1 | tail -6
for g in tc tc-bf16; do
timeout 300 python bench.py --gemm $g --no-cpu-baseline --no-e2e --gae-sweep 4096 > gpurun_out/bench_$g.json 2> gpurun_out/bench_$g.err
python -c "
import json
d=json.loads(open('gpurun_out/bench_$g.json').read().strip().splitlines()[-1])
print('$g', d['value'], d['ms_per_step'], d['roofline']['frac'], d['kernel_ms_per_step'], d['clocks'])
"
tail -2 gpurun_out/bench_$g.err
done
