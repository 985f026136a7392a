#!/bin/bash
mkdir -p gpurun_out

( time timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "off_the_256" ) > gpurun_out/c_pytest_off256.txt 2>&1; echo "off256 rc=$?"; tail -4 gpurun_out/c_pytest_off256.txt
( time timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 ) > gpurun_out/c_pytest_all.txt 2>&1; echo "all rc=$?"; tail -4 gpurun_out/c_pytest_all.txt
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" --encoder-dims 128,64 --decoder-dims 32 > gpurun_out/c_bench_128_64.json 2> gpurun_out/c_bench_128_64.err; echo "bench 128,64 rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" --encoder-dims 128,384,128 --decoder-dims 384 > gpurun_out/c_bench_384.json 2> gpurun_out/c_bench_384.err; echo "bench 384 rc=$?"
timeout 300 python bench.py --gemm fma --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" --encoder-dims 128,384,128 --decoder-dims 384 > gpurun_out/c_bench_384_fma.json 2> gpurun_out/c_bench_384_fma.err; echo "bench 384 fma rc=$?"
python - <<PY
import json
for f in ("c_bench_128_64", "c_bench_384", "c_bench_384_fma"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, "ms_per_step %.2f" % d["ms_per_step"], "value %.3g" % d["value"], {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()})
    except Exception as e:
        print(f, "failed", e)
PY
