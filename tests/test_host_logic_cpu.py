"""Host-side logic of the learner and the data-parallel layer (CPU, gloo)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ref_ppo
from oc_cleanrl_b200 import distributed as ocd
from oc_cleanrl_b200.learner import LearnerArgs, make_permutations


def test_args_arithmetic_matches_script():
    a = LearnerArgs(num_envs=4096, num_steps=128)
    assert (a.batch_size, a.minibatch_size) == (524288, 131072)
    o = ref_ppo.PPOArgs(num_envs=4096, num_steps=128)
    assert (o.batch_size, o.minibatch_size) == (a.batch_size, a.minibatch_size)
    d = LearnerArgs()
    assert (d.num_envs, d.clip_coef, d.ent_coef, d.vf_coef, d.max_grad_norm, d.learning_rate) == \
        (10, 0.1, 0.01, 0.5, 0.5, 2.5e-4)             # defaults of ppo_atari_oc.py, not ppo_atari.py


def test_permutations_follow_reference_sequence():
    rs1, rs2 = np.random.RandomState(42), np.random.RandomState(42)
    perms = make_permutations(1024, 4, rng=rs1, pinned=False).numpy()
    b = np.arange(1024)
    for e in range(4):
        rs2.shuffle(b)                                 # cumulative, ppo_atari_oc.py:558-561
        assert np.array_equal(perms[e], b)
    np.random.seed(3)
    p = make_permutations(64, 2, pinned=False).numpy()  # global numpy RNG, like the script
    np.random.seed(3)
    b = np.arange(64)
    np.random.shuffle(b)
    assert np.array_equal(p[0], b)


def test_shard_sizes_match_multigpu_script():
    for world in (1, 2, 4, 8):
        got = ocd.shard_sizes(512, 128, 4, world)
        assert got == ref_ppo.dp_world_math(512, 128, 4, world)
    assert ocd.shard_sizes(512, 128, 4, 8)["batch_size"] == 4096 * 128


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _dp_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    info = ocd.world_from_env()
    ocd.init_process_group(info, backend="gloo")
    rank_seed = ocd.seed_rank(1, info.rank)
    agent = ref_ppo.OraclePPObj(ref_ppo.stub_envs(4, 6, 6), "cpu", (16, 8), (8,))   # identical init on every rank
    ocd.reseed_sampling(rank_seed)
    perm = np.arange(32)
    np.random.shuffle(perm)                                   # differs per rank (seed + rank)
    flat = ref_ppo.flat_params(agent).clone()
    grad = torch.full_like(flat, float(rank + 1))
    ocd.allreduce_sum_(grad)
    out.put((rank, flat.double().sum().item(), perm[:8].tolist(), grad[0].item(), torch.rand(1).item()))
    dist.destroy_process_group()


def test_two_rank_gloo_seeding_and_flat_allreduce():
    """World-size-2 run of the DP host logic, as tests/test_atari_multigpu.py does for the
    reference (torchrun --nproc_per_node=2, gloo, CPU)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, w0, perm0, g0, u0), (r1, w1, perm1, g1, u1) = res
    assert w0 == w1                      # same initial weights without a broadcast (:211,:230)
    assert perm0 != perm1                # own minibatch permutation per rank (:208-210)
    assert g0 == g1 == 3.0               # SUM over ranks of the flat gradient (:367)
    assert u0 != u1                      # sampling stream re-seeded per rank (:231)


def test_dp_oracle_equals_single_rank_on_concatenated_batch_when_no_adv_norm():
    """Sanity of the K-rank simulation: with norm_adv off and equal local minibatches,
    mean-of-rank-gradients == gradient of the mean loss over the union."""
    torch.set_num_threads(1)
    args = ref_ppo.PPOArgs(num_envs=4, num_steps=8, norm_adv=False)
    mk = lambda: ref_ppo.make_agent(4, 6, 6, (16,), (8,), seed=3)  # noqa: E731
    g = torch.Generator().manual_seed(0)
    n = 16
    def batch():
        return (torch.randn(n, 4, 6, generator=g), torch.randint(0, 6, (n,), generator=g),
                -torch.rand(n, generator=g) - 1.0, torch.randn(n, generator=g), torch.randn(n, generator=g),
                torch.randn(n, generator=g))
    b0, b1 = batch(), batch()
    agents = [mk(), mk()]
    opts = [ref_ppo.make_optimizer(a, args) for a in agents]
    inds = [np.arange(n), np.arange(n)]
    _, gns, avg = ref_ppo.ref_dp_minibatch_step(agents, opts, [b0, b1], inds, args)
    single = mk()
    both = tuple(torch.cat([x, y]) for x, y in zip(b0, b1))
    loss, _ = ref_ppo.ref_minibatch_loss(single, *both, np.arange(2 * n), args)
    loss.backward()
    want = ref_ppo.flat_grads(single)
    assert torch.allclose(avg, want, rtol=1e-4, atol=1e-6)
    assert torch.equal(ref_ppo.flat_params(agents[0]), ref_ppo.flat_params(agents[1]))
    assert gns[0] == gns[1]
