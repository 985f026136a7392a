#!/bin/bash
mkdir -p gpurun_out
( time python tools/preact_err.py --engines 0,6,8 ) > gpurun_out/b_preact.txt 2>&1; echo "preact rc=$?"; tail -3 gpurun_out/b_preact.txt
python tools/gae_target.py 2 > gpurun_out/b_gae_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gae_scan -c 6 -o gpurun_out/b_gae_prof python tools/gae_target.py 2 > gpurun_out/b_gae_ncu.log 2>&1; echo "ncu rc=$?"
python tools/gae_bench.py > gpurun_out/b_gae_bench.txt 2>&1; cat gpurun_out/b_gae_bench.txt
python tools/rollout_bench.py > gpurun_out/b_rollout_bench.txt 2>&1; cat gpurun_out/b_rollout_bench.txt
python -m pytest tests/test_gpu_rollout.py -m gpu -q 2>&1 | tail -3
# traffic experiments on a bring-up build: what bounds the conversion-free mainloop?
OCPPO_BRINGUP=1 python -c "import __graft_entry__ as g; g.build(force=True)" > gpurun_out/b_build.log 2>&1
for d in 0 512 1024 1536; do
  OCPPO_BRINGUP=1 OCPPO_TC_DBG=$d timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e --no-configs --gae-sweep "" > gpurun_out/b_dbg$d.json 2> gpurun_out/b_dbg$d.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/b_dbg$d.json"))
    print("dbg $d: ms_per_step %.1f" % d["ms_per_step"], {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
except Exception as e:
    print("dbg $d failed", e)
PY
done
