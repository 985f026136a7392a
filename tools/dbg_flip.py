"""Which ReLU decisions of a 2,048-sample minibatch are within fp32 rounding of zero, and how much of the
tensor-core engine's gradient deviation from the CPU oracle sits in those units' rows (bring-up tool)."""
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
mb = 2048
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
perm = torch.as_tensor(np.random.RandomState(42).permutation(n)[:mb], dtype=torch.int64)
oracle = ref_ppo.make_agent(F, D, A, enc, dec, seed=42)
zs = []
hooks = [m.register_forward_hook(lambda m, i, o: zs.append((m, i[0].detach(), o.detach())))
         for m in oracle.network if isinstance(m, torch.nn.Linear)]
loss, _ = ref_ppo.ref_minibatch_loss(oracle, *cpu, perm.numpy(), ref_ppo.PPOArgs())
loss.backward()
for h in hooks: h.remove()
params = list(oracle.parameters())
want = [p.grad.reshape(-1).numpy().astype(np.float64) for p in params]
flag_rows = {}
for li, (m, x, z) in enumerate(zs):
    x, z = x.reshape(-1, x.shape[-1]), z.reshape(-1, z.shape[-1])
    absdot = x.abs().double() @ m.weight.abs().double().T + m.bias.abs().double()
    ratio = (z.double().abs() / absdot)
    for thr in (1e-7, 3e-7, 1e-6, 3e-6):
        print(f"layer {li} {tuple(m.weight.shape)}: |z| < {thr:g} * sum|w||x| : {(ratio < thr).sum().item()} decisions", flush=True)
    s, j = torch.nonzero(ratio < 1e-6, as_tuple=True)
    flag_rows[li] = sorted(set(j.tolist()))
    print("   flagged units", flag_rows[li])
assert lib.ocppo_set_gemm_mode(1) == 0
torch.manual_seed(42)
agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, enc, dec).to(dev)
learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
gpu = tuple(t.to(dev) for t in cpu)
flat = (gpu[0].reshape(n, -1),) + gpu[1:]
agent._ensure_workspace(mb, training=True)
_abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(perm.to(dev)), mb,
                                       C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
torch.cuda.synchronize()
got = [agent._grad_view(i).reshape(-1).double().cpu().numpy() for i, _ in enumerate(agent.parameters())]
shapes = [tuple(p.shape) for p in params]
tot_d2 = tot_w2 = ex_d2 = 0.0
for ti, (g, w) in enumerate(zip(got, want)):
    d = g - w
    shp = shapes[ti]
    rows = shp[0]
    d2 = (d.reshape(rows, -1) ** 2).sum(1)
    top = np.argsort(-d2)[:3]
    li = ti // 2
    fl = flag_rows.get(li, []) if ti < 2 * len(zs) else []
    keep = np.ones(rows, bool); keep[fl] = False
    print(f"tensor {ti} {shp}: L2 rel {np.sqrt(d2.sum() / (w ** 2).sum()):.2e}  maxnorm rel {np.abs(d).max() / np.abs(w).max():.2e}  "
          f"top rows {[(int(t), float(np.sqrt(d2[t] / (w ** 2).sum()))) for t in top]}  without flagged rows: L2 {np.sqrt(d2[keep].sum() / (w ** 2).sum()):.2e} "
          f"max {np.abs(d.reshape(rows, -1)[keep]).max() / np.abs(w).max():.2e}", flush=True)
    tot_d2 += d2.sum(); tot_w2 += (w ** 2).sum(); ex_d2 += d2[keep].sum()
print(f"whole gradient: L2 rel {np.sqrt(tot_d2 / tot_w2):.2e}; without flagged rows {np.sqrt(ex_d2 / tot_w2):.2e}")
