#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x > gpurun_out/m_pytest_gemm.txt 2>&1; echo "gemm tests rc=$?"; tail -1 gpurun_out/m_pytest_gemm.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "golden or full_size_minibatch_properties or off_the_256 or other_games or seeds_and_games" > gpurun_out/m_pytest_parity.txt 2>&1; echo "parity rc=$?"; tail -1 gpurun_out/m_pytest_parity.txt
timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/m_bench.json"))
print("ms_per_step %.2f" % d["ms_per_step"], {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items() if k.startswith("gemm")}, d["clocks"], "loss", d["last_stats"]["loss"])
PY
python tools/gemm_target.py 2 > gpurun_out/m_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -k regex:tc_gemm -s 12 -c 12 --csv --log-file gpurun_out/m_layers.csv python tools/gemm_target.py 2 > gpurun_out/m_ncu.log 2>&1
echo "per-launch ns (F1 F2 F3 F4 W4 D4 W3 D3 W2 D2 W1 D1):"; grep gpu__time_duration gpurun_out/m_layers.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%s ", $NF} END{print ""}'
echo "per-launch cycles:"; grep sm__cycles_elapsed gpurun_out/m_layers.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%s ", $NF} END{print ""}'
