#!/bin/bash
# Round-end validation on one B200: GPU tests, default bench, GEMM shape table, ncu launch list, one ncu --set full capture.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 --timeout-method thread > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/v_pytest.log
timeout 600 python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$?"; cat gpurun_out/v_bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/v_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --gae-sweep 4096 > gpurun_out/v_ncu_bench.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 6 -c 3 -f -o gpurun_out/v_tc8_full \
  python tools/gemm_bench.py 8 1 131072 > gpurun_out/v_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
