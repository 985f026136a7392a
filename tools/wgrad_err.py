"""WGRAD accumulation error at the trunk's real contraction lengths, per input kind (bring-up tool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gemm_diag
eng = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for kind in ("randn", "relu"):
    for (M, N, K, S) in ((1024, 512, 524288, 9), (512, 2048, 131072, 9), (512, 256, 524288, 32)):
        print(gemm_diag.run(2, eng, M, N, K, S, kind)[:160], flush=True)
