"""R12 on real hardware: the emulated two-rank test of tests/test_gpu_parity.py with the SUM done by NCCL over NVLink.
torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_nccl_parity.py
Every rank: one minibatch forward/backward on its own env shard (ocppo_minibatch_fwd_bwd), ncclAllReduce(SUM) in place on
the flat gradient arena, identical clip+Adam with 1/world folded in.  Checks: grad-norm and post-step weights against the
in-process K-rank oracle (ppo_atari_multigpu.py:360-377 restated, oracle/ref_ppo.py), weights bitwise equal across ranks.
Trunk widths are tensor-core sized so the default GEMM engine is the one exercised."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi  # noqa: E402
from oc_cleanrl_b200.architectures.ppo import PPObj  # noqa: E402
from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner  # noqa: E402
from oc_cleanrl_b200.synthetic import make_rollout  # noqa: E402
from oracle import ref_ppo  # noqa: E402  (checker only)

REL_TOL = 1e-4


def rel_err(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    F, D, A, T, N = 4, 6, 6, 32, 64
    enc, dec = (256, 512), (256,)
    n = T * N
    torch.manual_seed(1)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, enc, dec).to(dev)
    rs = np.random.RandomState(0)
    batches, idxs = [], []
    for r in range(world):                       # every rank builds all shards (the oracle on rank 0 needs them)
        data = make_rollout(T, N, F, D, rank=r)
        batches.append((data["obs"].reshape(n, F, D), torch.as_tensor(rs.randint(0, A, n)),
                        torch.as_tensor(-1.8 + 0.2 * rs.randn(n), dtype=torch.float32),
                        torch.as_tensor(rs.randn(n), dtype=torch.float32),
                        torch.as_tensor(rs.randn(n), dtype=torch.float32),
                        torch.as_tensor(rs.randn(n), dtype=torch.float32)))
        idxs.append(np.random.RandomState(100 + r).permutation(n)[: n // 4])
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T), world_size=1)
    lib = _abi.lib()
    flat = tuple(t.to(dev) for t in batches[rank])
    flat = (flat[0].reshape(n, -1),) + flat[1:]
    idx = torch.as_tensor(idxs[rank], dtype=torch.int64).to(dev)
    agent._ensure_workspace(idx.numel(), training=True)
    _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(idx), idx.numel(),
                                           C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
    dist.all_reduce(agent.grad_arena, op=dist.ReduceOp.SUM)            # ppo_atari_multigpu.py:362-368, in place
    gn = torch.zeros(1, device=dev)
    _abi.check(lib.ocppo_clip_adam(agent._handle.raw, 1.0 / world, 0.5, 2.5e-4, 0.9, 0.999, 1e-5, 1, _abi.ptr(gn),
                                   agent._stream()))
    mine = torch.cat([p.detach().reshape(-1) for p in agent.parameters()])
    everyone = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(everyone, mine)
    ok = True
    if rank == 0:
        same = all(torch.equal(everyone[0], e) for e in everyone[1:])
        oracles = [ref_ppo.make_agent(F, D, A, enc, dec, seed=1) for _ in range(world)]
        oargs = ref_ppo.PPOArgs(num_envs=N, num_steps=T)
        opts = [ref_ppo.make_optimizer(o, oargs) for o in oracles]
        _, gns, _ = ref_ppo.ref_dp_minibatch_step(oracles, opts, batches, idxs, oargs)
        e_gn = abs(float(gn) - float(gns[0])) / abs(float(gns[0]))
        e_w = rel_err(mine.cpu().numpy(), ref_ppo.flat_params(oracles[0]).numpy())
        ok = same and e_gn < REL_TOL and e_w < REL_TOL
        print(f"dp_nccl_parity world={world} engine={lib.ocppo_get_gemm_mode()} ranks_bitwise_equal={same} "
              f"grad_norm_rel_err={e_gn:.2e} weights_rel_err={e_w:.2e} -> {'PASS' if ok else 'FAIL'}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
