#!/bin/bash
# round-2 evidence pass on one B200: bench lines, accuracy tables, micro-benchmarks, then the ncu passes
mkdir -p gpurun_out
( time timeout 600 python bench.py --steps 10 --warmup 5 > gpurun_out/h_bench.json ) 2> gpurun_out/h_bench.err; echo "bench rc=$?"
( time timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/h_bench_ref.json ) 2> gpurun_out/h_bench_ref.err; echo "ref rc=$?"
( time timeout 900 python tools/fullsize_sweep.py --engines 8,6 --masks ) > gpurun_out/h_fullsize.txt 2>&1; echo "fullsize rc=$?"; tail -2 gpurun_out/h_fullsize.txt
python tools/preact_err.py --engines 0,6,8 > gpurun_out/h_preact.txt 2>&1; echo "preact rc=$?"
python tools/chained_update_err.py 0 6 8 > gpurun_out/h_chained.txt 2>&1; cat gpurun_out/h_chained.txt
python tools/small_mb_plain_err.py 0 6 8 > gpurun_out/h_small_mb.txt 2>&1; cat gpurun_out/h_small_mb.txt | cut -c1-200
python tools/gae_bench.py > gpurun_out/h_gae_bench.txt 2>&1; cat gpurun_out/h_gae_bench.txt
python tools/rollout_bench.py > gpurun_out/h_rollout_bench.txt 2>&1; cat gpurun_out/h_rollout_bench.txt
for cfg in "128,64 32" "128,384,128 384"; do set -- $cfg; tag=$(echo $1 | tr , _);
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" --encoder-dims $1 --decoder-dims $2 > gpurun_out/h_bench_enc_$tag.json 2> gpurun_out/h_bench_enc_$tag.err; echo "bench enc $1 rc=$?"
  timeout 300 python bench.py --gemm fma --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" --encoder-dims $1 --decoder-dims $2 > gpurun_out/h_bench_enc_${tag}_fma.json 2> gpurun_out/h_bench_enc_${tag}_fma.err; echo "bench enc $1 fma rc=$?"
done
# ncu: launch list of one plain-launch step, full capture of the second full-size minibatch's twelve GEMMs, GAE at the sweep sizes
CMD="python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep 4096"
$CMD > gpurun_out/h_plain.json 2> gpurun_out/h_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/h_launches.csv $CMD > gpurun_out/h_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python tools/gemm_target.py 2 > gpurun_out/h_gemm_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 12 -c 12 -o gpurun_out/h_gemm_prof python tools/gemm_target.py 2 > gpurun_out/h_gemm_ncu.log 2>&1; echo "ncu gemm rc=$?"
python tools/gae_target.py 1 > gpurun_out/h_gae_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gae_ -c 3 -o gpurun_out/h_gae_prof python tools/gae_target.py 1 > gpurun_out/h_gae_ncu.log 2>&1; echo "ncu gae rc=$?"
python tools/gemm_target.py 2 > gpurun_out/h_head_plain.log 2>&1 && ncu --set full --clock-control none -k regex:"loss_head|loss_bwd" -s 2 -c 2 -o gpurun_out/h_head_prof python tools/gemm_target.py 2 > gpurun_out/h_head_ncu.log 2>&1; echo "ncu head rc=$?"
ls -la gpurun_out | head -50
