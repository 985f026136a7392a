#!/bin/bash
# round-2 validation pass A: smoke, GPU tests, bench (b200 + reference arm), full-size error sweep
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/a_smi.txt 2>&1
( time python __graft_entry__.py --smoke ) > gpurun_out/a_smoke.txt 2>&1; echo "smoke rc=$?" 
( time timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 -x -k "graphs or seeds_and_games or rollout" ) > gpurun_out/a_pytest_new.txt 2>&1; echo "pytest(new) rc=$?"; tail -5 gpurun_out/a_pytest_new.txt
( time timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/a_bench.json ) 2> gpurun_out/a_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/a_bench.err
( time timeout 900 python tools/fullsize_sweep.py --engines 8,6 --masks ) > gpurun_out/a_fullsize.txt 2>&1; echo "fullsize rc=$?"; tail -3 gpurun_out/a_fullsize.txt
( time timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/a_bench_ref.json ) 2> gpurun_out/a_bench_ref.err; echo "ref rc=$?"
( time timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 -k "not graphs and not seeds_and_games and not rollout" ) > gpurun_out/a_pytest_old.txt 2>&1; echo "pytest(old) rc=$?"; tail -5 gpurun_out/a_pytest_old.txt
