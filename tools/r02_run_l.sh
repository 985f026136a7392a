#!/bin/bash
# lean MMA issue path: correctness (release build), per-launch times, then the main-product-only experiment on a bring-up build
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x > gpurun_out/l_pytest_gemm.txt 2>&1; echo "gemm tests rc=$?"; tail -1 gpurun_out/l_pytest_gemm.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "golden or full_size_minibatch_properties or off_the_256 or other_games or seeds_and_games" > gpurun_out/l_pytest_parity.txt 2>&1; echo "parity rc=$?"; tail -1 gpurun_out/l_pytest_parity.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/l_bench.json"))
print("release: ms_per_step %.2f" % d["ms_per_step"], {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items() if k.startswith("gemm")}, d["clocks"]["sm_mhz"], "loss", d["last_stats"]["loss"])
PY
python tools/gemm_target.py 2 > gpurun_out/l_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -k regex:tc_gemm -s 12 -c 12 --csv --log-file gpurun_out/l_layers.csv python tools/gemm_target.py 2 > gpurun_out/l_ncu.log 2>&1
echo "release per-launch ns (F1 F2 F3 F4 W4 D4 W3 D3 W2 D2 W1 D1):"; grep gpu__time_duration gpurun_out/l_layers.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%s ", $NF} END{print ""}'
echo "release per-launch cycles:"; grep sm__cycles_elapsed gpurun_out/l_layers.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%s ", $NF} END{print ""}'
OCPPO_BRINGUP=1 python -c "import __graft_entry__ as g; g.build(force=True)" > gpurun_out/l_build.log 2>&1; echo "bringup build rc=$?"
export OCPPO_BRINGUP=1
for d in 32 1568; do
  OCPPO_TC_DBG=$d ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -k regex:tc_gemm -s 12 -c 12 --csv --log-file gpurun_out/l_layers_dbg$d.csv python tools/gemm_target.py 2 > gpurun_out/l_ncu_dbg$d.log 2>&1
  echo "dbg $d per-launch cycles:"; grep sm__cycles_elapsed gpurun_out/l_layers_dbg$d.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%s ", $NF} END{print ""}'
done
