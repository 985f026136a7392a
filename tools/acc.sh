#!/bin/bash
# per-GEMM rounding error of the engines in $1.. on trunk-like shapes (tools/gemm_diag.py), randn and post-ReLU operands
python - "$@" <<'PY'
import sys
sys.path.insert(0, "tools")
import gemm_diag
for e in [int(x) for x in sys.argv[1:]]:
    for (op, M, N, K, S, kind) in [(0, 16384, 512, 1024, 1, "randn"), (0, 16384, 512, 2048, 1, "relu"), (1, 16384, 1024, 512, 1, "randn"),
                                   (2, 512, 1024, 65536, 9, "randn"), (2, 512, 1024, 65536, 9, "relu"), (2, 1024, 512, 131072, 9, "tiny+relu")]:
        print(gemm_diag.run(op, e, M, N, K, S, kind, seed=3)[:120], flush=True)
PY
