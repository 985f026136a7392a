#!/bin/bash
# usage: tools/gpu_retry.sh <timeout> <gpus> <cmd>   -- gpurun, retried while the pod has no free slot (nothing is charged for those)
T=$1; G=$2; shift 2
for i in $(seq 1 40); do
  if [ "$G" = "1" ]; then OUT=$(/usr/local/graft/bin/gpurun --timeout $T -- "$@" 2>&1); else OUT=$(/usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$@" 2>&1); fi
  if echo "$OUT" | grep -qE "status=transient|answers busy|rc=3"; then echo "[retry $i] no slot"; sleep 75; continue; fi
  echo "$OUT" | tail -70; exit 0
done
echo "gave up"
