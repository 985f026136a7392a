#!/bin/bash
# round-2 closing evidence pass on one B200 (after the loss-head / first-layer / skinny kernel rewrites): GPU suite, the
# driver's default bench line, accuracy tables, launch list, ncu --set full of the rewritten CUDA-core kernels
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/x_pytest_all.txt 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/x_pytest_all.txt | tail -2
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/x_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/x_smoke.txt
( time timeout 600 python bench.py ) > gpurun_out/x_bench.json 2> gpurun_out/x_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/x_bench.err
( time timeout 900 python tools/fullsize_sweep.py --engines 8 --masks ) > gpurun_out/x_fullsize.txt 2>&1; echo "fullsize rc=$?"; tail -2 gpurun_out/x_fullsize.txt
python tools/chained_update_err.py 0 8 > gpurun_out/x_chained.txt 2>&1; cat gpurun_out/x_chained.txt
python tools/small_mb_plain_err.py 0 8 > gpurun_out/x_small_mb.txt 2>&1; cat gpurun_out/x_small_mb.txt | cut -c1-200
python tools/rollout_bench.py > gpurun_out/x_rollout_bench.txt 2>&1; cat gpurun_out/x_rollout_bench.txt
timeout 300 python bench.py --num-envs 512 --steps 10 --warmup 5 --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/x_bench_512.json 2> gpurun_out/x_bench_512.err; echo "512 rc=$?"
ARGS='--steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep 4096'
python bench.py $ARGS > gpurun_out/x_plain.json 2> gpurun_out/x_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/x_launches.csv python bench.py $ARGS > gpurun_out/x_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"skinny_wgrad|thin_fwd|loss_head_quad|loss_bwd_pair" -s 8 -c 8 -o gpurun_out/x_small_prof -f python bench.py $ARGS > gpurun_out/x_ncu_small.log 2>&1; echo "ncu small rc=$?"
ls -la gpurun_out | grep " x_"
