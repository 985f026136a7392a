"""Device-side check of the raw GEMM entry point (both engines) against torch fp64.
Usage: python tools/gemm_diag.py [engine ...]      (default: 0 1)
Prints one line per case: max relative error and, on failure, where the errors sit."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi  # noqa: E402

lib = _abi.lib()
dev = torch.device("cuda:0")
P = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731


def run(op, engine, M, N, K, splits=1, kind="randn", seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    def mk(*shape):
        if kind == "int":
            return torch.randint(-4, 5, shape, device=dev, generator=g).float()
        if kind == "tiny":
            return torch.randn(*shape, device=dev, generator=g) * 1e-7
        if kind == "wide":          # nine decades of dynamic range inside one tensor
            e = torch.rand(*shape, device=dev, generator=g) * 9.0 - 6.0
            return torch.randn(*shape, device=dev, generator=g) * torch.pow(10.0, e)
        if kind == "big":
            return torch.randn(*shape, device=dev, generator=g) * 1e15
        if kind == "small":
            return torch.randn(*shape, device=dev, generator=g) * 1e-15
        if kind == "zero":
            return torch.zeros(*shape, device=dev)
        if kind == "relu":          # post-ReLU-like: half exact zeros, large positive values
            return torch.relu(torch.randn(*shape, device=dev, generator=g)) * 50.0
        return torch.randn(*shape, device=dev, generator=g)
    bias = mask = bias_part = None
    kp = -(-K // splits)
    kp = (kp + 31) // 32 * 32
    splits = -(-K // kp)
    kindB = None
    if "+" in kind:
        kind, kindB = kind.split("+")
    if op == 0:
        A, B = mk(M, K), mk(N, K)
        bias = mk(N)
        ref = torch.relu(A.double() @ B.double().t() + bias.double())
        C_ = torch.full((M, N), float("nan"), device=dev)
    elif op == 1:
        A, B = mk(M, K), mk(K, N)
        mask = mk(M, N)
        ref = (A.double() @ B.double()) * (mask > 0)
        C_ = torch.full((M, N), float("nan"), device=dev)
    else:
        A = mk(K, M)
        if kindB:
            kind = kindB
        B = mk(K, N)
        ref = A.double().t() @ B.double()
        C_ = torch.full((splits, M, N), float("nan"), device=dev)
        bias_part = torch.full((splits, M), float("nan"), device=dev)
    lda, ldb = A.shape[1], B.shape[1]
    rc = lib.ocppo_gemm(op, engine, M, N, K, P(A), lda, P(B), ldb, P(C_), N, P(bias), P(mask), P(bias_part),
                        splits, kp, None)
    if rc != 0:
        return f"rc={rc} {lib.ocppo_last_error().decode()}"
    torch.cuda.synchronize()
    out = C_.double().sum(0) if op == 2 else C_.double()
    err = (out - ref).abs()
    scale = ref.abs().max().item() or 1.0
    rel = err.max().item() / scale
    msg = f"op={op} eng={engine} M={M} N={N} K={K} S={splits} {kind}: rel_err={rel:.3e}"
    if op == 2:
        bref = A.double().sum(0)
        brel = ((bias_part.double().sum(0) - bref).abs().max() / (bref.abs().max() + 1e-30)).item()
        msg += f" bias_rel={brel:.3e}"
    if not (rel < 2e-5) or torch.isnan(out).any():
        bad = (err > 1e-4 * scale) | torch.isnan(out)
        idx = bad.nonzero()
        msg += f"  BAD {int(bad.sum())}/{bad.numel()} nan={int(torch.isnan(out).sum())}"
        if len(idx):
            r, c = idx[:, 0].cpu().numpy(), idx[:, 1].cpu().numpy()
            msg += (f" rows[{r.min()}..{r.max()}] cols[{c.min()}..{c.max()}] row%8={np.bincount(r % 8, minlength=8).tolist()}"
                    f" col%32/4={np.bincount((c % 32) // 4, minlength=8).tolist()} first={idx[:4].tolist()}"
                    f" got={out[idx[0, 0], idx[0, 1]].item():.4f} want={ref[idx[0, 0], idx[0, 1]].item():.4f}")
    return msg


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "probe":
    for e in (0, 1, 2):
        for (op, M, N, K, S, kind) in [(2, 1024, 512, 8192, 9, "tiny+relu"), (2, 1024, 512, 4096, 9, "tiny+relu"),
                                       (2, 512, 1024, 8192, 9, "tiny+relu"), (2, 512, 256, 8192, 32, "tiny+relu"),
                                       (2, 512, 2048, 2048, 4, "tiny+relu"), (1, 8192, 1024, 512, 1, "tiny"),
                                       (1, 2048, 2048, 512, 1, "tiny"), (1, 8192, 512, 1024, 1, "tiny")]:
            print(run(op, e, M, N, K, S, kind)[:150], flush=True)
    sys.exit(0)
if __name__ == "__main__":
    engines = [int(x) for x in sys.argv[1:]] or [0, 1]
    cases = [
        (0, 128, 256, 32, 1, "int"), (0, 128, 256, 64, 1, "int"), (0, 128, 256, 32, 1, "randn"),
        (0, 256, 512, 256, 1, "randn"), (0, 4096, 512, 256, 1, "randn"), (0, 1000, 256, 96, 1, "randn"),
        (1, 128, 256, 32, 1, "int"), (1, 128, 256, 64, 1, "int"), (1, 512, 256, 512, 1, "randn"),
        (1, 4096, 1024, 512, 1, "randn"),
        (2, 128, 256, 32, 1, "int"), (2, 128, 256, 64, 1, "int"), (2, 256, 256, 4096, 4, "randn"),
        (2, 512, 1024, 8192, 8, "randn"), (2, 256, 512, 1000, 2, "randn"),
        # persistent-kernel coverage: many more tiles than SMs, ragged M, several tiles per CTA
        (0, 32768, 1024, 512, 1, "randn"), (0, 20000, 512, 256, 1, "randn"), (1, 32768, 512, 1024, 1, "randn"),
        (1, 19999, 2048, 512, 1, "randn"), (2, 1024, 512, 65536, 9, "randn"), (2, 512, 2048, 32768, 19, "randn"),
    ]
    for e in engines:
        for (op, M, N, K, S, kind) in cases:
            print(run(op, e, M, N, K, S, kind), flush=True)
