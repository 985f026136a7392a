"""GAE scan timing at the BASELINE sweep sizes (rotating >L2 worth of buffers, CUDA events): python tools/gae_bench.py [iters]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

dev = torch.device("cuda:0")
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
pk = bench.peaks()
for n in (4096, 16384, 65536):
    us, us_copy, us_pdl = bench.time_gae_isolated(dev, 128, n, iters=iters)
    gbs = bench.gae_bytes(128, n) / (us * 1e-6) / 1e9
    print(f"GAE T=128 N={n}: {us:.2f} us/launch  {gbs:.0f} GB/s  {100 * gbs / pk['hbm']:.1f}% of {pk['hbm']:.0f} GB/s ({pk['source']}); "
          f"same-bytes copy {us_copy:.2f} us; as programmatic dependent launches {us_pdl:.2f} us = "
          f"{100 * bench.gae_bytes(128, n) / (us_pdl * 1e-6) / 1e9 / pk['hbm']:.1f}%", flush=True)


