#!/bin/bash
# A/B of the epilogue staging policy (OCPPO_TC_EPI2, bring-up build): correctness, per-layer kernel time, step time
mkdir -p gpurun_out
OCPPO_BRINGUP=1 python -c "import __graft_entry__ as g; g.build(force=True)" > gpurun_out/i_build.log 2>&1; echo "bringup build rc=$?"
export OCPPO_BRINGUP=1
for pol in 3 2; do
  OCPPO_TC_EPI2=$pol timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x > gpurun_out/i_pytest_gemm_pol$pol.txt 2>&1; echo "gemm tests policy $pol rc=$?"; tail -1 gpurun_out/i_pytest_gemm_pol$pol.txt
done
OCPPO_TC_EPI2=2 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "golden or full_size_minibatch_properties or off_the_256 or other_games" > gpurun_out/i_pytest_parity_pol2.txt 2>&1; echo "parity policy 2 rc=$?"; tail -1 gpurun_out/i_pytest_parity_pol2.txt
for pol in 0 1 2 3; do
  OCPPO_TC_EPI2=$pol timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/i_bench_pol$pol.json 2> gpurun_out/i_bench_pol$pol.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/i_bench_pol$pol.json"))
    print("policy $pol: ms_per_step %.2f" % d["ms_per_step"], {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items() if k.startswith("gemm")}, d["clocks"]["sm_mhz"], "loss", d["last_stats"]["loss"])
except Exception as e:
    print("policy $pol failed", e)
PY
done
python tools/gemm_target.py 2 > gpurun_out/i_plain.log 2>&1 || exit 1
for pol in 0 2 3; do
  OCPPO_TC_EPI2=$pol ncu --metrics gpu__time_duration.sum --clock-control none -k regex:tc_gemm -s 12 -c 12 --csv --log-file gpurun_out/i_layers_pol$pol.csv python tools/gemm_target.py 2 > gpurun_out/i_ncu_pol$pol.log 2>&1
  echo "policy $pol per-launch us (F1 F2 F3 F4 W4 D4 W3 D3 W2 D2 W1 D1):"; grep gpu__time_duration gpurun_out/i_layers_pol$pol.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%s ", $NF} END{print ""}'
done
