"""Where do ReLU decisions differ?  One full-size minibatch forward (131,072 samples, default widths): per trunk layer and
GEMM engine, the error of the post-ReLU activations against torch fp64 in units of the dot product's magnitude sum
S = |x| |W|^T + |b| (rms and max over elements positive on both sides), and the number of ReLU decisions that differ from
fp64's -- next to the same figures for torch fp32 (cuBLAS SGEMM, TF32 off).
Usage: python tools/preact_err.py [--game pong] [--seed 0] [--engines 0,6,8]"""
import argparse, ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
from oc_cleanrl_b200.synthetic import GAME_SHAPES, make_rollout, stub_envs
from oracle import ref_ppo
from fullsize import F, T, N, ENC, DEC, MB

p = argparse.ArgumentParser()
p.add_argument("--game", default="pong"); p.add_argument("--seed", type=int, default=0); p.add_argument("--engines", default="0,6,8")
a = p.parse_args()
dev = torch.device("cuda:0")
lib = _abi.lib()
D, A = GAME_SHAPES[a.game]
n = T * N
seed = a.seed
cpu_model = ref_ppo.make_agent(F, D, A, ENC, DEC, seed=42 + seed)
data = make_rollout(T, N, F, D, seed=1234 + 17 * seed)
rs = np.random.RandomState(1 + seed)
gpu = (data["obs"].reshape(n, F, D).to(dev), torch.as_tensor(rs.randint(0, A, n)).to(dev),
       torch.as_tensor(-np.log(A) + 0.05 * rs.randn(n), dtype=torch.float32).to(dev),
       torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev), torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev),
       torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev))
flat = (gpu[0].reshape(n, -1),) + gpu[1:]
perm = torch.as_tensor(np.random.RandomState(42 + seed).permutation(n)[:MB], dtype=torch.int64).to(dev)
x0 = gpu[0][perm]

def torch_acts(dtype):
    o = ref_ppo.make_agent(F, D, A, ENC, DEC, seed=42 + seed, dtype=dtype).to(dev)
    acts, h = [], x0.to(dtype)
    with torch.no_grad():
        for m in o.network:
            h = m(h)
            if isinstance(m, torch.nn.ReLU):
                acts.append(h.reshape(-1, h.shape[-1]))
    return o, acts

o64, a64 = torch_acts(torch.float64)
lins = [m for m in o64.network if isinstance(m, torch.nn.Linear)]
S = []
with torch.no_grad():
    h = x0.double().reshape(-1, D)
    for l, lin in enumerate(lins):
        if l == len(ENC):
            h = h.reshape(MB, -1)
        S.append(h.abs() @ lin.weight.abs().T + lin.bias.abs())
        h = a64[l]

def report(tag, acts):
    for l, (got, want, s) in enumerate(zip(acts, a64, S)):
        got = got.double()
        both = (got > 0) & (want > 0)
        e = ((got - want) / s)[both]
        flips = int(((got > 0) != (want > 0)).sum())
        print(f"{tag} layer {l} {tuple(want.shape)}: rel-to-magnitude-sum err rms {float(e.pow(2).mean().sqrt()):.2e} max {float(e.abs().max()):.2e}  "
              f"mean {float(e.mean()):+.2e}  decisions differing from fp64: {flips} of {want.numel()}", flush=True)
        del got, both, e

_, a32 = torch_acts(torch.float32)
report("torch fp32", a32)
del a32
for engine in [int(x) for x in a.engines.split(",")]:
    assert lib.ocppo_set_gemm_mode(engine) == 0
    torch.manual_seed(42 + seed)
    agent = PPObj(stub_envs(F, D, A), dev, ENC, DEC).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
    agent._ensure_workspace(MB, training=True, exact=True)
    _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(perm), MB,
                                           C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
    acts = []
    for l, width in enumerate(ENC + DEC):
        act = torch.empty(MB * (F if l < len(ENC) else 1), width, device=dev)
        _abi.check(lib.ocppo_read_activation(agent._handle.raw, l, MB, _abi.ptr(act), agent._stream()))
        acts.append(act)
    torch.cuda.synchronize()
    report(f"engine {engine}", acts)
    del acts, agent, learner
    torch.cuda.empty_cache()
