"""A 2,048-sample minibatch at default widths: per-tensor gradient error against the CPU oracle WITHOUT handing it the
device path's ReLU decisions (the plain comparison), per GEMM engine, next to the number of differing decisions.  At this
size one flipped decision is worth 1/2048 of a gradient row, so the plain comparison can exceed 1e-4 on the tensor-core
engines although every GEMM is fp32-grade (tests/test_gpu_parity.py compares under the device path's
decisions and proves each differing one undecidable).   python tools/small_mb_plain_err.py [engines...]"""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
from oc_cleanrl_b200.synthetic import make_rollout, stub_envs
from oracle import ref_ppo

dev = torch.device("cuda:0")
lib = _abi.lib()
F, D, A, T, N, mb = 4, 6, 6, 128, 4096, 2048
enc, dec = (256, 512, 1024, 512), (512,)
data = make_rollout(T, N, F, D)
n = T * N
rs = np.random.RandomState(1)
cpu = (data["obs"].reshape(n, F, D), torch.as_tensor(rs.randint(0, A, n)),
       torch.as_tensor(-1.79 + 0.05 * rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32), torch.as_tensor(rs.randn(n), dtype=torch.float32),
       torch.as_tensor(rs.randn(n), dtype=torch.float32))
gpu = tuple(t.to(dev) for t in cpu)
flat = (gpu[0].reshape(n, -1),) + gpu[1:]
perm = np.random.RandomState(42).permutation(n)[:mb]
idx = torch.as_tensor(perm, dtype=torch.int64).to(dev)
oracle = ref_ppo.make_agent(F, D, A, enc, dec, seed=42)
with ref_ppo.relu_decisions(oracle) as rd:
    loss, _ = ref_ppo.ref_minibatch_loss(oracle, *cpu, perm, ref_ppo.PPOArgs())
loss.backward()
want = [p.grad.reshape(-1).double().numpy() for p in oracle.parameters()]
own = [r["decision"].reshape(-1, r["decision"].shape[-1]) for r in rd.records]
rel = lambda a, b: float(np.abs(a - b).max() / np.abs(b).max())
for engine in [int(x) for x in sys.argv[1:]] or [0, 6, 8]:
    assert lib.ocppo_set_gemm_mode(engine) == 0
    torch.manual_seed(42)
    agent = PPObj(stub_envs(F, D, A), dev, enc, dec).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
    agent._ensure_workspace(mb, training=True, exact=True)
    _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(idx), mb,
                                           C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
    torch.cuda.synchronize()
    got = [agent._grad_view(i).reshape(-1).double().cpu().numpy() for i in range(len(want))]
    flips = 0
    for l, width in enumerate(enc + dec):
        act = torch.empty(mb * (F if l < len(enc) else 1), width, device=dev)
        _abi.check(lib.ocppo_read_activation(agent._handle.raw, l, mb, _abi.ptr(act), agent._stream()))
        torch.cuda.synchronize()
        flips += int(((act > 0).cpu() != own[l]).sum())
    errs = [rel(a, b) for a, b in zip(got, want)]
    print(f"engine {engine}: plain comparison vs the CPU oracle, worst tensor {max(errs):.2e}  [{' '.join(f'{e:.1e}' for e in errs)}]  "
          f"ReLU decisions differing from the oracle's own: {flips} of {sum(o.numel() for o in own)}", flush=True)
