"""Feeds REAL layer operands (captured from a torch fp32 run of the reference restatement) to the raw GEMM
entry point and compares every engine with fp64 (bring-up tool)."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.synthetic import make_rollout
from oracle import ref_ppo

dev = torch.device("cuda:0")
lib = _abi.lib()
P = lambda t: None if t is None else C.c_void_p(t.data_ptr())
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
F, D, A, T, N = 4, 6, 6, 128, 64
enc, dec = (256, 512, 1024, 512), (512,)
data = make_rollout(T, N, F, D)
n = T * N
rs = np.random.RandomState(1)
cpu = (data["obs"].reshape(n, F, D), torch.as_tensor(rs.randint(0, A, n)),
       torch.as_tensor(-1.79 + 0.05 * rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32), torch.as_tensor(rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32))
perm = np.random.RandomState(42).permutation(n)[:mb]
oracle = ref_ppo.make_agent(F, D, A, enc, dec, seed=42).to(dev)
gpu = tuple(t.to(dev) for t in cpu)
acts, grads = {}, {}
def fwd_hook(name):
    def f(mod, inp, out):
        acts[name] = inp[0].detach()
        out.register_hook(lambda g: grads.__setitem__(name, g.detach()))
    return f
for i in (2, 4, 6):
    oracle.network[i].register_forward_hook(fwd_hook(i))
loss, _ = ref_ppo.ref_minibatch_loss(oracle, *gpu, torch.as_tensor(perm, device=dev), ref_ppo.PPOArgs())
loss.backward()
for i in (2, 4, 6):
    X = acts[i].reshape(-1, acts[i].shape[-1]).contiguous()        # [rows, in]
    dY = grads[i].reshape(-1, grads[i].shape[-1]).contiguous()      # [rows, out]  (grad wrt pre-activation)
    K, Mo, Ni = X.shape[0], dY.shape[1], X.shape[1]
    ref = dY.double().t() @ X.double()
    t32 = (dY.t() @ X).double()
    canc = float((dY.double().abs().t() @ X.double().abs()).max() / ref.abs().max())
    msg = f"layer {i}: dW[{Mo},{Ni}] K={K} cancellation {canc:.1f}  torch32 {float((t32-ref).abs().max()/ref.abs().max()):.2e}"
    for eng in (0, 1, 2):
        for S in (9, 32):
            kp = (-(-K // S) + 31) // 32 * 32
            Sx = -(-K // kp)
            Cc = torch.zeros(Sx, Mo, Ni, device=dev)
            bp = torch.zeros(Sx, Mo, device=dev)
            rc = lib.ocppo_gemm(2, eng, Mo, Ni, K, P(dY), Mo, P(X), Ni, P(Cc), Ni, None, None, P(bp), Sx, kp, None)
            assert rc == 0, lib.ocppo_last_error().decode()
            torch.cuda.synchronize()
            e = float((Cc.double().sum(0) - ref).abs().max() / ref.abs().max())
            msg += f" | e{eng}/S{Sx} {e:.2e}"
    print(msg, flush=True)
    print("    |dY| max %.2e rms %.2e  |X| max %.2e mean %.2e  zeros in X %.2f" % (
        float(dY.abs().max()), float(dY.pow(2).mean().sqrt()), float(X.abs().max()), float(X.mean()), float((X == 0).float().mean())))
