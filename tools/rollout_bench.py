"""Rollout-step plumbing at the benchmark size (4096 envs, Pong-shaped, default widths): steps per second of the
reference's own per-step data flow (ppo_atari_oc.py:500-514, pageable torch.tensor uploads, action.cpu()) against the
RolloutCollector (pinned staging, in-place kernel outputs, CUDA-graph replay), with a host environment that costs
nothing -- and the host env feed: the per-step HOST time of the reference's per-env Python observation
assembly + VecNormalize reward path (oracle/ref_feed.py restatement) against ocppo_feed_push / _normalize_reward.
Usage: python tools/rollout_bench.py [num_envs=4096] [steps=128] [slots=3]"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.rollout import RolloutCollector
from oc_cleanrl_b200.feed import ObjectFeed
from oc_cleanrl_b200.synthetic import stub_envs

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = int(sys.argv[2]) if len(sys.argv) > 2 else 128
S = int(sys.argv[3]) if len(sys.argv) > 3 else 3
F, D, A = 4, 2 * S, 6
dev = torch.device("cuda:0")
torch.manual_seed(42)
agent = PPObj(stub_envs(F, D, A), dev, (256, 512, 1024, 512), (512,)).to(dev)
rs = np.random.RandomState(0)
pool = [(rs.randint(0, 200, (N, F, D)).astype(np.float64), rs.randn(N), rs.rand(N) < 0.005) for _ in range(8)]

def reference_style():
    obs = torch.zeros((T, N, F, D), device=dev); actions = torch.zeros((T, N), dtype=torch.long, device=dev)
    logprobs = torch.zeros((T, N), device=dev); rewards = torch.zeros((T, N), device=dev)
    dones = torch.zeros((T, N), device=dev); values = torch.zeros((T, N), device=dev)
    next_obs = torch.tensor(pool[0][0], dtype=torch.float32, device=dev); next_done = torch.zeros(N, device=dev)
    def run():
        nonlocal next_obs, next_done
        for step in range(T):
            obs[step] = next_obs
            dones[step] = next_done
            with torch.no_grad():
                action, logprob, _, value = agent.get_action_and_value(next_obs)
                values[step] = value.flatten()
            actions[step] = action
            logprobs[step] = logprob
            _ = action.detach().cpu().numpy()
            o, r, d = pool[step % 8]
            rewards[step] = torch.tensor(r, dtype=torch.float32, device=dev).view(-1)
            next_obs = torch.tensor(o, dtype=torch.float32, device=dev)
            next_done = torch.tensor(d, dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
    return run

def collector_style(graphs=True):
    col = RolloutCollector(agent, T, N, use_graphs=graphs)
    col.reset(pool[0][0])
    def run():
        for step in range(T):
            col.act(step)
            col.observe(step, *pool[step % 8])
        torch.cuda.synchronize()
    return run

slot_pool = [(np.stack([rs.randint(0, 160, (N, S)), rs.randint(0, 210, (N, S))], -1).astype(np.int32), (rs.rand(N, S) > 0.1),
              (rs.rand(N) < 0.1) * 1.0, rs.rand(N) < 0.005) for _ in range(8)]

def collector_feed_style():
    col = RolloutCollector(agent, T, N, feed=ObjectFeed(N, S, F))
    col.reset_objects(slot_pool[0][0], slot_pool[0][1])
    def run():
        for step in range(T):
            col.act(step)
            col.observe_objects(step, *slot_pool[step % 8])
        torch.cuda.synchronize()
    return run

for name, make in (("reference data flow", reference_style), ("RolloutCollector, plain launches", lambda: collector_style(False)),
                   ("RolloutCollector, graph replay", collector_style), ("RolloutCollector + object feed", collector_feed_style)):
    run = make()
    run()
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        run()
    dt = (time.perf_counter() - t0) / reps
    print(f"{name:>34s}: {dt / T * 1e3:.3f} ms per env step of {N} envs, {N * T / dt / 1e6:.2f} M env-steps/s (host env cost excluded)", flush=True)

# ---- host side alone: observation assembly + reward normalisation per env step of N envs ----
from oracle import ref_feed  # noqa: E402  (the reference's per-env Python path, restated; timing baseline only)
n_ref = min(N, 256)                       # the Python path is per env: time a slice and scale
envs = [ref_feed.RefObjectEnv(S, F) for _ in range(n_ref)]
norm = ref_feed.RefVecNormalizeReward(N)
for n, e in enumerate(envs):
    e.reset(slot_pool[0][0][n], slot_pool[0][1][n])
t0 = time.perf_counter()
reps = 20
for i in range(reps):
    xy, vis, reward, done = slot_pool[i % 8]
    obs = np.stack([e.step(xy[n], vis[n]) for n, e in enumerate(envs)])
    norm.step_wait(reward.copy(), done)
t_ref = (time.perf_counter() - t0) / reps * (N / n_ref)
feed = ObjectFeed(N, S, F)
out = np.empty((N, F, D), np.float32)
feed.push(slot_pool[0][0], slot_pool[0][1], out=out)
t0 = time.perf_counter()
reps = 200
for i in range(reps):
    xy, vis, reward, done = slot_pool[i % 8]
    feed.push(xy, vis, done, out=out)
    feed.normalize_reward(reward, done)
t_feed = (time.perf_counter() - t0) / reps
print(f"host assembly + reward normalisation per env step of {N} envs x {S} slots: reference-style Python path {t_ref * 1e3:.2f} ms "
      f"(in the reference spread over the env subprocesses), ocppo_feed {t_feed * 1e3:.3f} ms", flush=True)
