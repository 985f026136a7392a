"""Per-tensor gradient error of ONE full-size minibatch (131,072 samples of a 4096 x 128 Pong-shaped rollout, default
widths) against the same operator sequence run by torch on the device in fp32 and fp64, per GEMM engine.
Usage: python tools/fullsize_err.py [modes...]   (default 0 1 3)"""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
from oc_cleanrl_b200.synthetic import make_rollout
from oracle import ref_ppo

dev = torch.device("cuda:0")
F, D, A, T, N = 4, 6, 6, 128, 4096
enc, dec = (256, 512, 1024, 512), (512,)
lib = _abi.lib()
data = make_rollout(T, N, F, D)
n = T * N
rs = np.random.RandomState(1)
cpu = (data["obs"].reshape(n, F, D), torch.as_tensor(rs.randint(0, A, n)),
       torch.as_tensor(-1.79 + 0.05 * rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32), torch.as_tensor(rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32))
gpu = tuple(t.to(dev) for t in cpu)
flat = (gpu[0].reshape(n, -1),) + gpu[1:]
perm = torch.as_tensor(np.random.RandomState(42).permutation(n), dtype=torch.int64).to(dev)
mb = 131072

def torch_grads(dtype, masks=None):
    o = ref_ppo.make_agent(F, D, A, enc, dec, seed=42, dtype=dtype).to(dev)
    args_t = tuple(t.to(dtype) if t.is_floating_point() else t for t in gpu)
    if masks is None:
        l, _ = ref_ppo.ref_minibatch_loss(o, *args_t, perm[:mb], ref_ppo.PPOArgs())
    else:
        with ref_ppo.relu_decisions(o, masks):
            l, _ = ref_ppo.ref_minibatch_loss(o, *args_t, perm[:mb], ref_ppo.PPOArgs())
    l.backward()
    g = [q.grad.reshape(-1).double().cpu().numpy() for q in o.parameters()]
    del o, args_t, l
    torch.cuda.empty_cache()
    return g

g32, g64 = torch_grads(torch.float32), torch_grads(torch.float64)
rel = lambda a, b: float(np.abs(a - b).max() / np.abs(b).max())
print("torch fp32 vs fp64: " + " ".join(f"{rel(a, b):.1e}" for a, b in zip(g32, g64)), flush=True)
for mode in [int(x) for x in sys.argv[1:]] or [0, 1, 3]:
    assert lib.ocppo_set_gemm_mode(mode) == 0
    torch.manual_seed(42)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, enc, dec).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
    agent._ensure_workspace(mb, training=True)
    _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(perm), mb,
                                           C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
    torch.cuda.synchronize()
    got = [agent._grad_view(i).reshape(-1).double().cpu().numpy() for i, _ in enumerate(agent.parameters())]
    print(f"mode {mode} vs fp32: " + " ".join(f"{rel(a, b):.1e}" for a, b in zip(got, g32)), flush=True)
    print(f"mode {mode} vs fp64: " + " ".join(f"{rel(a, b):.1e}" for a, b in zip(got, g64)), flush=True)
    masks = []
    for l, width in enumerate(enc + dec):
        act = torch.empty(mb * (F if l < len(enc) else 1), width, device=dev)
        _abi.check(lib.ocppo_read_activation(agent._handle.raw, l, mb, _abi.ptr(act), agent._stream()))
        torch.cuda.synchronize()
        masks.append(act > 0)
        del act
    g64m = torch_grads(torch.float64, masks)
    print(f"mode {mode} vs fp64 under the device path's ReLU decisions: " + " ".join(f"{rel(a, b):.1e}" for a, b in zip(got, g64m)), flush=True)
    del masks, g64m
    del agent, learner
    torch.cuda.empty_cache()
