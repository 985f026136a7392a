"""Trunk forward output (hidden) per GEMM engine vs the oracle, row by row (bring-up tool)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.synthetic import make_rollout
from oracle import ref_ppo

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
F, D, A = 4, 6, 6
enc, dec = (256, 512, 1024, 512), (512,)
lib = _abi.lib()
x = make_rollout(128, 64, F, D)["obs"].reshape(-1, F, D)[:B].contiguous()
oracle = ref_ppo.make_agent(F, D, A, enc, dec, seed=42).double()
with torch.no_grad():
    want = oracle.network(x.double()).numpy()
for mode in (0, 1, 2):
    assert lib.ocppo_set_gemm_mode(mode) == 0
    torch.manual_seed(42)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, enc, dec).to(dev)
    agent._ensure_workspace(B, training=True)
    hidden = torch.empty(B, 512, device=dev)
    _abi.check(lib.ocppo_trunk_forward(agent._handle.raw, _abi.ptr(x.to(dev)), B, _abi.ptr(hidden), agent._stream()))
    torch.cuda.synchronize()
    got = hidden.double().cpu().numpy()
    err = np.abs(got - want)
    scale = np.abs(want).max()
    rowerr = err.max(1) / scale
    bad = np.nonzero(rowerr > 2e-5)[0]
    flips = int(((got > 0) != (want > 0)).sum())
    print(f"mode {mode}: max rel err {err.max()/scale:.2e}  median row err {np.median(rowerr):.2e}  rows>2e-5: {len(bad)}"
          f"  first {bad[:12].tolist()}  mask flips {flips}/{got.size}", flush=True)
