"""Per-tensor gradient error of one minibatch vs the CPU oracle, per GEMM engine (bring-up tool)."""
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
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
F, D, A, T, N = 4, 6, 6, 128, 64
enc, dec = (256, 512, 1024, 512), (512,)
lib = _abi.lib()
data = make_rollout(T, N, F, D)
n = T * N
rs = np.random.RandomState(1)
cpu = (data["obs"].reshape(n, F, D), torch.as_tensor(rs.randint(0, A, n)),
       torch.as_tensor(-1.79 + 0.05 * rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32), torch.as_tensor(rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32))
perm = torch.as_tensor(np.random.RandomState(42).permutation(n)[:mb], dtype=torch.int64)
oracle = ref_ppo.make_agent(F, D, A, enc, dec, seed=42)
loss, _ = ref_ppo.ref_minibatch_loss(oracle, *cpu, perm.numpy(), ref_ppo.PPOArgs())
loss.backward()
want = [p.grad.reshape(-1).numpy().astype(np.float64) for p in oracle.parameters()]
for mode in (1, 2):
    assert lib.ocppo_set_gemm_mode(mode) == 0
    torch.manual_seed(42)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, enc, dec).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
    gpu = tuple(t.to(dev) for t in cpu)
    flat = (gpu[0].reshape(n, -1),) + gpu[1:]
    agent._ensure_workspace(mb, training=True)
    for rep in range(1):
        _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(perm.to(dev)), mb,
                                               C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
        torch.cuda.synchronize()
        got = [agent._grad_view(i).reshape(-1).double().cpu().numpy() for i, _ in enumerate(agent.parameters())]
        errs = [float(np.abs(g - w).max() / np.abs(w).max()) for g, w in zip(got, want)]
        print(f"mode {mode} rep {rep}: " + " ".join(f"{e:.1e}" for e in errs), flush=True)
        shapes = [tuple(p.shape) for p in agent.parameters()]
        for ti in (5, 4, 9):
            d = np.abs(got[ti] - want[ti]) / np.abs(want[ti]).max()
            top = np.argsort(-d)[:6]
            cols = shapes[ti][1] if len(shapes[ti]) == 2 else 1
            print(f"   tensor {ti} {shapes[ti]}: top errors " + ", ".join(f"({t // cols},{t % cols}):{d[t]:.1e}" for t in top)
                  + f"  | entries > 1e-5: {(d > 1e-5).sum()} of {d.size}, median {np.median(d):.1e}")
