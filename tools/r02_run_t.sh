#!/bin/bash
# ncu --set full of the CUDA-core kernels around the trunk (first-layer forward, skinny wgrads, loss head pair)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep ''"
python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep "" > gpurun_out/t_plain.json 2> gpurun_out/t_plain.err && ncu --set full --clock-control none --import-source on -k regex:"skinny_wgrad|thin_fwd|loss_head_quad|loss_bwd_pair" -s 8 -c 8 -o gpurun_out/t_small_prof -f python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep "" > gpurun_out/t_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/t_small_prof.ncu-rep
