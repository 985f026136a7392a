"""Rollout-step plumbing: the collector against the reference's own per-step data flow
(cleanrl/ppo_atari_oc.py:500-514) driven through the public agent API, on a deterministic host environment."""
import numpy as np
import pytest
import torch

from oracle import ref_ppo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev(built_lib):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


class ToyVecEnv:
    """Deterministic stand-in for the SubprocVecEnv of the script: integer-valued object vectors, sparse rewards,
    occasional episode ends; everything a function of (step counter, actions)."""

    def __init__(self, n, frames, obs_dim, seed=0):
        self.n, self.frames, self.obs_dim, self.t = n, frames, obs_dim, 0
        self.state = np.random.RandomState(seed).randint(0, 160, (n, frames, obs_dim)).astype(np.float64)

    def reset(self):
        self.t = 0
        return self.state.copy()

    def step(self, actions):
        assert actions.dtype == np.int64 and actions.shape == (self.n,)
        self.t += 1
        self.state = np.roll(self.state, 1, axis=1)
        self.state[:, 0] = np.clip(self.state[:, 1] + (actions[:, None] - 2) + (self.t % 3), 0, 209)
        reward = np.where(actions == self.t % 6, 1.0, 0.0) - 0.25 * (actions == 0)
        done = (actions == (self.t * 7) % 6) & (np.arange(self.n) % 5 == self.t % 5)
        return self.state.copy(), reward, done, [{}] * self.n


def reference_style_rollout(agent, env, dev, T, N, iterations, seed):
    """ppo_atari_oc.py:451-466 and :500-514, statement for statement, with the drop-in agent."""
    F, D = agent._frames, agent._obs_dim
    obs = torch.zeros((T, N, F, D), dtype=torch.float32, device=dev)
    actions = torch.zeros((T, N), dtype=torch.long, device=dev)
    logprobs = torch.zeros((T, N), dtype=torch.float32, device=dev)
    rewards = torch.zeros((T, N), dtype=torch.float32, device=dev)
    dones = torch.zeros((T, N), dtype=torch.float32, device=dev)
    values = torch.zeros((T, N), dtype=torch.float32, device=dev)
    torch.cuda.manual_seed(seed)
    next_obs = torch.tensor(env.reset(), dtype=torch.float32, device=dev)
    next_done = torch.zeros(N, dtype=torch.float32, device=dev)
    snaps = []
    for _ in range(iterations):
        for step in range(T):
            obs[step] = next_obs
            dones[step] = next_done
            with torch.no_grad():
                action, logprob, _, value = agent.get_action_and_value(next_obs)
                values[step] = value.flatten()
            actions[step] = action
            logprobs[step] = logprob
            next_obs_np, reward_np, next_done_np, _ = env.step(action.detach().cpu().numpy())
            rewards[step] = torch.tensor(reward_np, dtype=torch.float32, device=dev).view(-1)
            next_obs = torch.tensor(next_obs_np, dtype=torch.float32, device=dev)
            next_done = torch.tensor(next_done_np, dtype=torch.float32, device=dev)
        snaps.append([t.clone() for t in (obs, actions, logprobs, rewards, dones, values, next_obs, next_done)])
    return snaps


@pytest.mark.parametrize("T,N,dims", [(16, 64, ((32, 64), (32,))), (8, 512, ((256, 512, 1024, 512), (512,)))])
def test_collector_reproduces_the_reference_data_flow_bit_for_bit(dev, T, N, dims):
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.rollout import RolloutCollector
    F, D, A = 4, 6, 6
    torch.manual_seed(7)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, *dims).to(dev)
    want = reference_style_rollout(agent, ToyVecEnv(N, F, D), dev, T, N, iterations=2, seed=11)

    env = ToyVecEnv(N, F, D)
    col = RolloutCollector(agent, T, N)
    torch.cuda.manual_seed(11)
    col.reset(env.reset())
    for it in range(2):
        if it:
            col.start_next_iteration()
        for step in range(T):
            a = col.act(step)
            assert a.dtype == np.int64 and a.shape == (N,)
            col.observe(step, *env.step(a.copy())[:3])
        torch.cuda.synchronize()
        got = (col.obs, col.actions, col.logprobs, col.rewards, col.dones, col.values, col.next_obs, col.next_done)
        for name, g, w in zip(("obs", "actions", "logprobs", "rewards", "dones", "values", "next_obs", "next_done"),
                              got, want[it]):
            assert g.shape == w.shape and g.dtype == w.dtype, name
            assert torch.equal(g, w), f"iteration {it}: {name} differs"
        assert col.obs.is_contiguous() and col.dones.is_contiguous()


def test_collector_feeds_the_learner_and_rejects_bad_input(dev):
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    from oc_cleanrl_b200.rollout import RolloutCollector
    F, D, A, T, N = 4, 6, 6, 8, 32
    torch.manual_seed(3)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, (32, 64), (32,)).to(dev)
    env = ToyVecEnv(N, F, D)
    col = RolloutCollector(agent, T, N)
    col.reset(env.reset())
    for step in range(T):
        col.observe(step, *env.step(col.act(step).copy())[:3])
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T), np_rng=np.random.RandomState(0))
    adv, ret = learner.compute_gae(col.rewards, col.values, col.dones, col.next_obs, col.next_done)
    stats = learner.update(col.obs, col.actions, col.logprobs, adv, ret, col.values)
    assert np.isfinite(stats["loss"]) and stats["minibatches"] == 16
    with pytest.raises(_abi.OcppoError):
        col.act(T)
    with pytest.raises(_abi.OcppoError):
        col.observe(0, np.zeros((N + 1, F, D)), np.zeros(N), np.zeros(N))


def test_collector_fed_with_object_slots_equals_the_reference_env_side(dev):
    """object slots in (what an env worker reads off the emulator), through ``ocppo_feed_push`` /
    ``ocppo_feed_normalize_reward`` into the pinned staging buffers, against the reference's env side restated in
    oracle/ref_feed.py (per-env list chain + deque, SB3 VecNormalize reward path) feeding the array entry points."""
    from oracle import ref_feed
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.feed import ObjectFeed
    from oc_cleanrl_b200.rollout import RolloutCollector
    F, S, A, T, N = 4, 3, 6, 12, 48
    D = 2 * S
    rng = np.random.RandomState(5)
    script = [(np.stack([rng.randint(0, 160, (N, S)), rng.randint(0, 210, (N, S))], -1), rng.rand(N, S) > 0.2,
               (rng.rand(N) < 0.1) * rng.choice([1.0, -1.0, 20.0], N), rng.rand(N) < 0.1) for _ in range(2 * T + 1)]
    bufs = []
    for use_feed in (False, True):
        torch.manual_seed(3)
        agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, (32, 64), (32,)).to(dev)
        torch.cuda.manual_seed(4)
        if use_feed:
            col = RolloutCollector(agent, T, N, feed=ObjectFeed(N, S, F))
            col.reset_objects(script[0][0], script[0][1])
        else:
            col = RolloutCollector(agent, T, N)
            envs = [ref_feed.RefObjectEnv(S, F) for _ in range(N)]
            norm = ref_feed.RefVecNormalizeReward(N)
            col.reset(np.stack([e.reset(script[0][0][n], script[0][1][n]) for n, e in enumerate(envs)]))
        k = 1
        for it in range(2):
            for step in range(T):
                col.act(step)
                xy, vis, reward, done = script[k]
                k += 1
                if use_feed:
                    col.observe_objects(step, xy, vis, reward, done)
                else:
                    obs = np.stack([e.reset(xy[n], vis[n]) if done[n] else e.step(xy[n], vis[n]) for n, e in enumerate(envs)])
                    col.observe(step, obs, norm.step_wait(reward.copy(), done), done)
        torch.cuda.synchronize()
        bufs.append([t.clone() for t in (col.obs, col.actions, col.logprobs, col.rewards, col.dones, col.values, col.next_obs, col.next_done)])
    for a, b in zip(*bufs):
        assert torch.equal(a, b)


@pytest.mark.parametrize("n,const", [(128 * 4096, False), (1000, False), (77, True)])
def test_explained_variance_matches_the_numpy_line(dev, n, const):
    """ppo_atari_oc.py:620-623 (np.var on host copies) against the device reduction."""
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    torch.manual_seed(1)
    agent = PPObj(ref_ppo.stub_envs(4, 6, 6), dev, (32, 64), (32,)).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=8, num_steps=8))
    rs = np.random.RandomState(n)
    y_true = np.full(n, 3.5, np.float32) if const else (2.0 + 1.5 * rs.randn(n)).astype(np.float32)
    y_pred = (y_true + 0.3 * rs.randn(n)).astype(np.float32)
    var_y = np.var(y_true)
    want = np.nan if var_y == 0 else 1 - np.var(y_true - y_pred) / var_y
    got = learner.explained_variance(torch.as_tensor(y_pred).to(dev), torch.as_tensor(y_true).to(dev))
    if const:
        assert np.isnan(want) and np.isnan(got)
    else:
        assert got == pytest.approx(float(want), rel=1e-4, abs=1e-6)


def test_checkpoint_round_trips_with_the_reference_format(dev, tmp_path):
    """ppo_atari_oc.py:484-491 / loading.py:111-122: a file written here loads into the reference-shaped
    module, a file written the reference's way loads here, and training resumes bit-identically."""
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.checkpoint import load_checkpoint, save_checkpoint
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    from oc_cleanrl_b200.synthetic import make_rollout
    F, D, A, T, N = 4, 6, 6, 8, 32
    dims = ((32, 64), (32,))
    torch.manual_seed(5)
    agent = PPObj(ref_ppo.stub_envs(F, D, A), dev, *dims).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T), np_rng=np.random.RandomState(1))
    data = make_rollout(T, N, F, D)
    rs = np.random.RandomState(2)
    roll = [data["obs"].to(dev), torch.as_tensor(rs.randint(0, A, (T, N))).to(dev),
            torch.as_tensor(-1.79 + 0.05 * rs.randn(T, N), dtype=torch.float32).to(dev),
            torch.as_tensor(rs.randn(T, N), dtype=torch.float32).to(dev),
            torch.as_tensor(rs.randn(T, N), dtype=torch.float32).to(dev),
            torch.as_tensor(rs.randn(T, N), dtype=torch.float32).to(dev)]
    learner.update(*roll)
    path = tmp_path / "run_1.cleanrl_model"
    saved = save_checkpoint(path, agent, args={"env_id": "ALE/Pong-v5", "seed": 1}, timesteps=T * N, learner=learner)
    assert list(saved["model_weights"].keys()) == list(agent.state_dict().keys()) and saved["Timesteps"] == T * N
    # the reference-shaped module reads it the way loading.py does
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    oracle = ref_ppo.make_agent(F, D, A, *dims, seed=0)
    oracle.load_state_dict(ckpt["model_weights"])
    x = data["obs"].reshape(-1, F, D)[:64]
    with torch.no_grad():
        assert torch.allclose(agent.get_value(x.to(dev)).cpu(), oracle.get_value(x), rtol=1e-4, atol=1e-5)
    # continue training, then restore into a fresh agent and continue again: identical weights
    perm_rng = np.random.RandomState(9)
    learner.np_rng = perm_rng
    learner.update(*roll)
    want = agent.arena.clone()
    torch.manual_seed(99)
    agent2 = PPObj(ref_ppo.stub_envs(F, D, A), dev, *dims).to(dev)
    learner2 = PPOLearner(agent2, LearnerArgs(num_envs=N, num_steps=T), np_rng=np.random.RandomState(9))
    load_checkpoint(path, agent2, learner2)
    assert learner2.adam_step == 16
    learner2.update(*roll)
    assert torch.equal(agent2.arena, want)
    # a file written the reference's way (weights only, :675-677)
    path2 = tmp_path / "final.cleanrl_model"
    torch.save({"model_weights": oracle.state_dict(), "args": {}}, path2)
    load_checkpoint(path2, agent2)
    assert all(torch.equal(a.cpu(), b) for a, b in zip(agent2.state_dict().values(), oracle.state_dict().values()))
