"""Times the raw GEMM entry point at the trunk's real shapes (Pong, 32768-sample chunk).
Usage: python tools/gemm_bench.py [engine=1] [iters=5]   -> one line per GEMM: ms, effective TFLOP/s."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi  # noqa: E402

lib = _abi.lib()
dev = torch.device("cuda:0")
P = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731


def bench(op, engine, M, N, K, splits, iters):
    g = torch.Generator(device=dev).manual_seed(0)
    mk = lambda *s: torch.randn(*s, device=dev, generator=g)  # noqa: E731
    kp = (-(-K // splits) + 31) // 32 * 32
    splits = -(-K // kp)
    bias = mask = bias_part = None
    if op == 0:
        A, B, bias, Cc = mk(M, K), mk(N, K), mk(N), torch.empty(M, N, device=dev)
    elif op == 1:
        A, B, mask, Cc = mk(M, K), mk(K, N), mk(M, N), torch.empty(M, N, device=dev)
    else:
        A, B, Cc = mk(K, M), mk(K, N), torch.empty(splits, M, N, device=dev)
        bias_part = torch.empty(splits, M, device=dev)
    def run():
        rc = lib.ocppo_gemm(op, engine, M, N, K, P(A), A.shape[1], P(B), B.shape[1], P(Cc), N, P(bias), P(mask),
                            P(bias_part), splits, kp, None)
        assert rc == 0, lib.ocppo_last_error().decode()
    for _ in range(2):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    tf = 2.0 * M * N * K / (ms * 1e-3) / 1e12
    print(f"op={op} eng={engine} M={M} N={N} K={K} S={splits}: {ms:.3f} ms  {tf:.1f} TFLOP/s (fp32-equivalent)", flush=True)


if __name__ == "__main__":
    engine = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    R = int(sys.argv[3]) if len(sys.argv) > 3 else 32768
    shapes = [(0, 4 * R, 512, 256, 1), (0, 4 * R, 1024, 512, 1), (0, 4 * R, 512, 1024, 1), (0, R, 512, 2048, 1),
              (1, 4 * R, 256, 512, 1), (1, 4 * R, 512, 1024, 1), (1, 4 * R, 1024, 512, 1), (1, R, 2048, 512, 1),
              (2, 512, 256, 4 * R, 32), (2, 1024, 512, 4 * R, 9), (2, 512, 1024, 4 * R, 9), (2, 512, 2048, R, 9)]
    for s in shapes:
        bench(s[0], engine, s[1], s[2], s[3], s[4], iters)
