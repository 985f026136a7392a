#!/bin/bash
# round-2 final evidence on one B200 (after the TMA-issuer and decision-word changes to tc_gemm): default bench line, launch list,
# ncu --set full of the second minibatch's twelve GEMMs
mkdir -p gpurun_out
( time timeout 600 python bench.py ) > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/fin_bench.err
( time timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/fin_bench_ref.json ) 2> gpurun_out/fin_bench_ref.err; echo "ref rc=$?"
python tools/chained_update_err.py 0 8 > gpurun_out/fin_chained.txt 2>&1; cat gpurun_out/fin_chained.txt
ARGS='--steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep 4096'
python bench.py $ARGS > gpurun_out/fin_plain.json 2> gpurun_out/fin_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/fin_launches.csv python bench.py $ARGS > gpurun_out/fin_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python tools/gemm_target.py 2 > gpurun_out/fin_gemm_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 12 -c 12 -f -o gpurun_out/fin_gemm_prof python tools/gemm_target.py 2 > gpurun_out/fin_gemm_ncu.log 2>&1; echo "ncu gemm rc=$?"
ls -la gpurun_out | grep " fin_"
