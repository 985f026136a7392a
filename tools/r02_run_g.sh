#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > gpurun_out/g_pytest_zz_alone.txt 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/g_pytest_zz_alone.txt
( time timeout 1800 python -m pytest tests -x -q -m gpu ) > gpurun_out/g_pytest_all.txt 2>&1; echo "all (-x) rc=$?"; tail -4 gpurun_out/g_pytest_all.txt
( time python __graft_entry__.py --smoke ) > gpurun_out/g_smoke.txt 2>&1; echo "smoke rc=$?"; grep smoke gpurun_out/g_smoke.txt
