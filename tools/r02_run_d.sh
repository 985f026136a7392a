#!/bin/bash
mkdir -p gpurun_out
python tools/gae_cases.py > gpurun_out/d_gae_cases.txt 2>&1; cat gpurun_out/d_gae_cases.txt
( time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "gae" ) > gpurun_out/d_pytest_gae.txt 2>&1; echo "gae tests rc=$?"; tail -4 gpurun_out/d_pytest_gae.txt
python tools/gae_bench.py > gpurun_out/d_gae_bench.txt 2>&1; cat gpurun_out/d_gae_bench.txt
( time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "off_the_256" ) > gpurun_out/d_pytest_off256.txt 2>&1; echo "off256 rc=$?"; tail -4 gpurun_out/d_pytest_off256.txt
( time timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 ) > gpurun_out/d_pytest_all.txt 2>&1; echo "all rc=$?"; tail -4 gpurun_out/d_pytest_all.txt
( time python __graft_entry__.py --smoke ) > gpurun_out/d_smoke.txt 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/d_smoke.txt
