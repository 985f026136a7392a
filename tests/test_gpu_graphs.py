"""CUDA-graph replay, launch accounting, device selection and the real 2-GPU NCCL step (needs a B200: ``-m gpu``).

* a replayed graph must leave exactly what the plain launches leave (bitwise), for the update and for the rollout step;
* the profiling hook must time every launch it counts (round 1 dropped launches past a fixed-size ring);
* an agent on a device that is not current must compute on ITS device (the reference's multi-GPU script never calls
  ``torch.cuda.set_device``, ppo_atari_multigpu.py:216-222);
* R12 on real hardware: ``tools/dp_nccl_parity.py`` under ``torch.distributed.run`` when two GPUs are visible.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import ref_ppo
from oc_cleanrl_b200.synthetic import make_rollout, stub_envs

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def dev(built_lib):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _agent(dev, F, D, A, enc, dec, seed=42):
    from oc_cleanrl_b200.architectures.ppo import PPObj
    torch.manual_seed(seed)
    return PPObj(stub_envs(F, D, A), dev, enc, dec).to(dev)


def _flat(agent):
    return torch.cat([p.detach().reshape(-1) for p in agent.parameters()])


@pytest.mark.parametrize("dims", [((256, 512), (256,)), ((64, 128), (64,))])
def test_update_graph_replay_equals_plain_launches_bitwise(dev, dims):
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    F, D, A, T, N = 4, 6, 6, 16, 64
    data = make_rollout(T, N, F, D, seed=5)
    results = []
    for use_graphs in (False, True):
        agent = _agent(dev, F, D, A, *dims)
        largs = LearnerArgs(num_envs=N, num_steps=T)
        learner = PPOLearner(agent, largs, np_rng=np.random.RandomState(3), use_graphs=use_graphs)
        d = {k: v.to(dev) for k, v in data.items()}
        torch.cuda.manual_seed(11)
        with torch.no_grad():
            action, logp, _, value = agent.get_action_and_value(d["obs"].reshape(T * N, F, D))
        stats = []
        for it in range(3):                      # update 0 captures, 1 and 2 replay; the LR changes every iteration (:471-474)
            learner.anneal_lr(it + 1, 10)
            adv, ret = learner.compute_gae(d["rewards"], value.reshape(T, N), d["dones"], d["next_obs"], d["next_done"])
            stats.append(learner.update(d["obs"], action.reshape(T, N), logp.reshape(T, N), adv, ret, value.reshape(T, N)))
        results.append((_flat(agent).clone(), agent.exp_avg.clone(), agent.exp_avg_sq.clone(), stats, learner))
    (w0, m0, v0, s0, _), (w1, m1, v1, s1, lg) = results
    assert lg.use_graphs and lg._graph_hits == 32 and lg._graph_misses == 16
    assert torch.equal(w0, w1) and torch.equal(m0, m1) and torch.equal(v0, v1)
    for a, b in zip(s0, s1):
        assert a == b


def test_rollout_step_graph_replay_equals_plain_launches_bitwise(dev):
    """Same torch CUDA seed, same observations: the graph-replayed rollout step must sample the same actions and consume
    the generator identically, call after call (the Philox offset lives in device memory and advances per call); after
    an optimizer step the replayed graph must see the new weights."""
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    F, D, A, B = 4, 6, 6, 512
    enc, dec = (256, 512), (256,)
    data = make_rollout(8, B, F, D, seed=9)
    obs = data["obs"].to(dev)                                  # (8, B, F, D): eight persistent "rows"
    outs = []
    for graph in (False, True):
        agent = _agent(dev, F, D, A, enc, dec)
        bufs = (torch.zeros(8, B, dtype=torch.int64, device=dev), torch.zeros(8, B, device=dev),
                torch.zeros(B, device=dev), torch.zeros(8, B, device=dev))
        torch.cuda.manual_seed(123)
        for rnd in range(3):
            for t in range(8):
                agent.policy_step_into(obs[t], bufs[0][t], bufs[1][t], bufs[2], bufs[3][t], graph=graph)
            if rnd == 1:                                       # weights change between rollouts
                learner = PPOLearner(agent, LearnerArgs(num_envs=B, num_steps=8), np_rng=np.random.RandomState(0), use_graphs=graph)
                adv, ret = learner.compute_gae(data["rewards"].to(dev), bufs[3], data["dones"].to(dev),
                                               data["next_obs"].to(dev), data["next_done"].to(dev))
                learner.update(obs, bufs[0], bufs[1], adv, ret, bufs[3])
        gen = torch.cuda.default_generators[dev.index]
        outs.append(([b.clone() for b in bufs], gen.get_offset(), _flat(agent).clone()))
    (b0, off0, w0), (b1, off1, w1) = outs
    assert off0 == off1
    assert torch.equal(w0, w1)
    for x, y in zip(b0, b1):
        assert torch.equal(x, y)


def test_profiling_times_every_launch_it_counts(dev, built_lib):
    """More launches than the event ring holds (2048): none may be dropped."""
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200.learner import compute_gae
    T, N = 4, 64
    r, v, d = (torch.rand(T, N, device=dev) for _ in range(3))
    nv, nd = torch.rand(N, device=dev), torch.zeros(N, device=dev)
    out = (torch.empty(T, N, device=dev), torch.empty(T, N, device=dev))
    c0, _ = _abi.launch_counts()
    built_lib.ocppo_profile_enable(1)
    for _ in range(5000):
        compute_gae(r, v, d, nv, nd, 0.99, 0.95, out=out)
    ms, timed = _abi.profile_read()
    built_lib.ocppo_profile_enable(0)
    c1, _ = _abi.launch_counts()
    assert timed["gae"] == 5000 == c1["gae"] - c0["gae"]
    assert sum(timed.values()) == 5000
    assert 0 < ms["gae"] < 5000 * 1.0


def test_graph_captured_launches_are_reported_untimed(dev, built_lib):
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200.learner import compute_gae
    T, N = 4, 64
    r, v, d = (torch.rand(T, N, device=dev) for _ in range(3))
    nv, nd = torch.rand(N, device=dev), torch.zeros(N, device=dev)
    out = (torch.empty(T, N, device=dev), torch.empty(T, N, device=dev))
    compute_gae(r, v, d, nv, nd, 0.99, 0.95, out=out)
    torch.cuda.synchronize()
    built_lib.ocppo_profile_enable(1)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        compute_gae(r, v, d, nv, nd, 0.99, 0.95, out=out)
    g.replay()
    torch.cuda.synchronize()
    _, timed = _abi.profile_read()
    _, untimed = _abi.launch_counts()
    built_lib.ocppo_profile_enable(0)
    assert timed["gae"] == 0 and untimed["gae"] == 1


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_agent_on_a_device_that_is_not_current(built_lib):
    """ppo_atari_multigpu.py:216-222 builds ``device = cuda:{local_rank}`` and never makes it current."""
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    F, D, A, T, N = 4, 6, 6, 8, 32
    enc, dec = (256, 512), (256,)
    data = make_rollout(T, N, F, D, seed=2)
    res = []
    for ordinal in (0, 1):
        torch.cuda.set_device(0)                          # current device stays 0 throughout
        dev = torch.device("cuda", ordinal)
        agent = _agent(dev, F, D, A, enc, dec)
        d = {k: v.to(dev) for k, v in data.items()}
        with torch.no_grad():
            value = agent.get_value(d["obs"].reshape(T * N, F, D))
            action, logp, _, _ = agent.get_action_and_value(d["obs"].reshape(T * N, F, D), torch.zeros(T * N, dtype=torch.int64, device=dev))
        learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T), np_rng=np.random.RandomState(1))
        adv, ret = learner.compute_gae(d["rewards"], value.reshape(T, N), d["dones"], d["next_obs"], d["next_done"])
        learner.update(d["obs"], action.reshape(T, N), logp.reshape(T, N), adv, ret, value.reshape(T, N))
        learner.update(d["obs"], action.reshape(T, N), logp.reshape(T, N), adv, ret, value.reshape(T, N))
        torch.cuda.synchronize(dev)
        assert torch.cuda.current_device() == 0
        res.append((value.cpu(), adv.cpu(), _flat(agent).cpu()))
    for a, b in zip(*res):
        assert torch.equal(a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_nccl_step_against_the_k_rank_oracle():
    """R12 on real hardware: two processes, NCCL over NVLink, against oracle/ref_ppo.py's K-rank restatement of
    ppo_atari_multigpu.py:360-377; mirrors how the reference tests its multi-GPU script (tests/test_atari_multigpu.py:4-9)."""
    port = 29000 + os.getpid() % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "dp_nccl_parity.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PASS" in r.stdout
