"""ncu target: a few GAE launches at the BASELINE sweep sizes over fresh (cold-L2) buffers.
python tools/gae_target.py [reps]   ->   launches are N = 4096, 16384, 65536, repeated `reps` times (default 2)"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200.learner import compute_gae  # noqa: E402

dev = torch.device("cuda:0")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
T = 128
g = torch.Generator(device=dev).manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for rep in range(reps):
    for N in (4096, 16384, 65536):
        r, v = torch.randn(T, N, device=dev, generator=g), torch.randn(T, N, device=dev, generator=g)
        d = (torch.rand(T, N, device=dev, generator=g) < 0.005).float()
        nv, nd = torch.randn(N, device=dev, generator=g), torch.zeros(N, device=dev)
        flush.zero_()                                   # evict the inputs from L2
        adv, ret = compute_gae(r, v, d, nv, nd, 0.99, 0.95)
        torch.cuda.synchronize()
        print(f"N={N}: adv checksum {float(adv.double().sum()):.6f}")
