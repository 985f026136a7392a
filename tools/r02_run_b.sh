#!/bin/bash
mkdir -p gpurun_out
( time python tools/preact_err.py --engines 0,6,8 ) > gpurun_out/b_preact.txt 2>&1; echo "preact rc=$?"; tail -3 gpurun_out/b_preact.txt
python tools/gae_target.py 2 > gpurun_out/b_gae_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gae_scan -c 6 -o gpurun_out/b_gae_prof python tools/gae_target.py 2 > gpurun_out/b_gae_ncu.log 2>&1; echo "ncu rc=$?"
python tools/gae_bench.py > gpurun_out/b_gae_bench.txt 2>&1; cat gpurun_out/b_gae_bench.txt
