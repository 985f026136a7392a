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
    us = bench.time_gae_isolated(dev, 128, n, iters=iters)
    gbs = bench.gae_bytes(128, n) / (us * 1e-6) / 1e9
    print(f"GAE T=128 N={n}: {us:.2f} us/launch  {gbs:.0f} GB/s  {100 * gbs / pk['hbm']:.1f}% of {pk['hbm']:.0f} GB/s ({pk['source']})", flush=True)


def time_copy_same_bytes(T, n, iters):
    """Yardstick: a device-to-device copy moving the same number of bytes per launch (2.5*T*N floats in, 2.5*T*N
    out), replayed from a graph over rotating >L2 buffers exactly like the GAE timing.  It shows how much of the
    large-buffer peak ANY kernel of this size reaches once launch and ramp-up are paid."""
    nel = (5 * T * n) // 2
    n_sets = max(2, int((192 << 20) // (nel * 8)) + 1)
    sets = [(torch.randn(nel, device=dev), torch.empty(nel, device=dev)) for _ in range(n_sets)]
    for s, d in sets:
        d.copy_(s)
    torch.cuda.synchronize(dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for i in range(iters):
            s, d = sets[i % n_sets]
            d.copy_(s)
    graph.replay()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) * 1e3 / iters


for n in (4096, 16384, 65536):
    us = time_copy_same_bytes(128, n, iters)
    gbs = bench.gae_bytes(128, n) / (us * 1e-6) / 1e9
    print(f"copy of the same bytes, N={n}: {us:.2f} us/launch  {gbs:.0f} GB/s  {100 * gbs / pk['hbm']:.1f}%", flush=True)
