"""Rollout-step plumbing at the benchmark size (4096 envs, Pong-shaped, default widths): steps per second of the
reference's own per-step data flow (ppo_atari_oc.py:500-514, pageable torch.tensor uploads, action.cpu()) against the
RolloutCollector (pinned staging, in-place kernel outputs), with a host environment that costs nothing.
Usage: python tools/rollout_bench.py [num_envs=4096] [steps=128]"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.rollout import RolloutCollector
from oracle.ref_ppo import stub_envs            # 3-attribute env stub only

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = int(sys.argv[2]) if len(sys.argv) > 2 else 128
F, D, A = 4, 6, 6
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

def collector_style():
    col = RolloutCollector(agent, T, N)
    col.reset(pool[0][0])
    def run():
        for step in range(T):
            col.act(step)
            col.observe(step, *pool[step % 8])
        torch.cuda.synchronize()
    return run

for name, make in (("reference data flow", reference_style), ("RolloutCollector", collector_style)):
    run = make()
    run()
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        run()
    dt = (time.perf_counter() - t0) / reps
    print(f"{name:>20s}: {dt / T * 1e3:.3f} ms per env step of {N} envs, {N * T / dt / 1e6:.2f} M env-steps/s (host env cost excluded)", flush=True)
