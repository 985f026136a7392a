#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --num-envs 512 --steps 10 --warmup 5 --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/p_bench_512.json 2> gpurun_out/p_bench_512.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/p_bench_512.json"))
print("512 envs: ms_per_step %.2f e2e %.2f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]), "profile pass %.2f sum kernels %.2f" % (d["profile_pass"]["ms_per_step"], d["profile_pass"]["sum_kernel_ms_per_step"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["profile_pass"]["kernels_per_step_by_class"], d["clocks"])
PY
CMD="python bench.py --num-envs 512 --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep ''"
python bench.py --num-envs 512 --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep "" > gpurun_out/p_plain.json 2> gpurun_out/p_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/p_launches.csv python bench.py --num-envs 512 --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep "" > gpurun_out/p_ncu.log 2>&1; echo "ncu rc=$?"
