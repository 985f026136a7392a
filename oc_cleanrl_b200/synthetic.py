"""Seeded synthetic object-centric rollouts.

Shapes follow OCAtari ``obs_mode="obj"``: an observation is ``(frames, D)`` with
``D = 2 * n_object_slots`` ([x, y] per slot, invisible objects are [0, 0]);
``frames`` stacked consecutive states.  Values are integer-valued fp32 pixel
coordinates, exactly what the reference feeds PPObj un-normalised
(architectures/ppo.py:89-90).  Generated on the host with a CPU generator so the
CPU oracle and the device path see the same bytes.
"""
from __future__ import annotations

from types import SimpleNamespace

import torch


def stub_envs(frames: int, obs_dim: int, n_actions: int):
    """A vector-env stand-in carrying the three attributes ``PPObj.__init__`` reads (architectures/ppo.py:65,83):
    ``observation_space.shape == (frames, obs_dim)`` and ``action_space.n``."""
    return SimpleNamespace(observation_space=SimpleNamespace(shape=(int(frames), int(obs_dim))),
                           action_space=SimpleNamespace(n=int(n_actions)))


GAME_SHAPES = {            # (obs_dim D, n_actions A): ocatari/ram/{pong,spaceinvaders,breakout}.py
    "pong": (6, 6),
    "spaceinvaders": (88, 6),
    "breakout": (220, 4),
}


def make_rollout(num_steps, num_envs, frames=4, obs_dim=6, seed=1234, rank=0, sparse_reward=True,
                 done_prob=1.0 / 200.0):
    """Returns a dict of CPU fp32 tensors: obs (T,N,F,D), rewards/dones (T,N),
    next_obs (N,F,D), next_done (N,)."""
    g = torch.Generator().manual_seed(seed + rank)
    T, N, F, D = num_steps, num_envs, frames, obs_dim
    hi = torch.empty(D)
    hi[0::2] = 159.0
    hi[1::2] = 209.0
    state = torch.floor(torch.rand((N, D), generator=g) * (hi + 1.0))
    visible = (torch.rand((N, (D + 1) // 2), generator=g) >= 0.10).float()      # (an odd width ends with half an object slot)
    visible = visible.repeat_interleave(2, dim=1)[:, :D]
    total = T + F                                 # states s_0 .. s_{T+F-1}
    states = torch.empty((total, N, D))
    for i in range(total):
        states[i] = state * visible
        step = torch.randint(-4, 5, (N, D), generator=g).float()
        state = torch.minimum(torch.clamp(state + step, min=0.0), hi)
    # obs[t, n, f] = state at (t + f): the newest frame is last, as a deque-based frame stack
    obs = torch.stack([states[f:f + T] for f in range(F)], dim=2).contiguous()
    next_obs = torch.stack([states[T + f] for f in range(F)], dim=1).contiguous()
    rewards = torch.randn((T, N), generator=g).clamp_(-10.0, 10.0)
    if sparse_reward:
        rewards = rewards * (torch.rand((T, N), generator=g) < 0.10).float()
    dones = (torch.rand((T, N), generator=g) < done_prob).float()
    next_done = (torch.rand((N,), generator=g) < done_prob).float()
    return dict(obs=obs, rewards=rewards, dones=dones, next_obs=next_obs, next_done=next_done)


def make_gae_inputs(num_steps, num_envs, seed=42):
    """Distributions of the reference's only numeric test
    (tests/test_jax_compute_gae.py:76-86): values~U(0,1), rewards~U(-1,1),
    dones~Bernoulli(0.5)."""
    g = torch.Generator().manual_seed(seed)
    T, N = num_steps, num_envs
    return dict(
        rewards=torch.rand((T, N), generator=g) * 2.0 - 1.0,
        values=torch.rand((T, N), generator=g),
        dones=(torch.rand((T, N), generator=g) < 0.5).float(),
        next_value=torch.rand((N,), generator=g),
        next_done=(torch.rand((N,), generator=g) < 0.5).float(),
    )
