"""Rollout-step plumbing around the fused inference kernel.

Mirrors the storage and the per-step data flow of ``cleanrl/ppo_atari_oc.py:451-466`` (buffers) and ``:500-514``
(one environment step), with three differences that leave every stored value identical:

* the reference builds three pageable ``torch.tensor(np_array, device=...)`` per step and reads the actions back with
  ``action.cpu().numpy()``; here the host side is four persistent PINNED staging buffers and the copies are
  asynchronous on the agent's stream;
* the reference copies ``next_obs`` / ``next_done`` into ``obs[step]`` / ``dones[step]`` on the device every step
  (``:502-503``); here the host arrays of step ``t`` are uploaded straight into row ``t + 1`` of the ``(T, N, ...)``
  buffers (row 0 of the next iteration via the carried ``next_obs`` / ``next_done``), so those copy kernels disappear;
* ``ocppo_policy_step`` takes raw pointers, so the fused kernel writes ``actions[step]``, ``logprobs[step]`` and
  ``values[step]`` in place (``:506-509`` are three more copies in the reference).

The only synchronisation per step is the one the environment needs: the actions of this step on the host.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _abi
from .architectures.ppo import PPObj


class RolloutCollector:
    """``(T, N, ...)`` device buffers exactly as the script allocates them plus pinned host staging.

    Typical use (the loop of ``ppo_atari_oc.py:500-514``)::

        col = RolloutCollector(agent, num_steps, num_envs)
        col.reset(envs.reset())                                   # :463-465
        for iteration in ...:
            for step in range(num_steps):
                actions = col.act(step)                           # :502-509 (+ the D2H of :511); act(0) first carries
                                                                  # next_obs / next_done of the previous iteration into row 0
                next_obs, reward, done, infos = envs.step(actions)
                col.observe(step, next_obs, reward, done)         # :512-514
            adv, ret = learner.compute_gae(col.rewards, col.values, col.dones, col.next_obs, col.next_done)
            learner.update(col.obs, col.actions, col.logprobs, adv, ret, col.values)
    """

    def __init__(self, agent: PPObj, num_steps: int, num_envs: int, use_graphs: bool = True, feed=None):
        """``use_graphs``: replay the launches of each rollout step from a CUDA graph (one per buffer row, captured on
        first use; see ``PPObj.policy_step_into``).  ``feed``: an :class:`oc_cleanrl_b200.feed.ObjectFeed` for
        ``reset_objects`` / ``observe_objects`` (object slots in, frame-stacked block written straight into the pinned
        staging buffer, rewards normalised like ``VecNormalize``)."""
        if agent._handle is None:
            raise _abi.OcppoError("move the agent to its CUDA device before building the collector")
        self.agent = agent
        self.T, self.N = int(num_steps), int(num_envs)
        self.device = dev = agent._cuda_device
        F, D = agent._frames, agent._obs_dim
        T, N = self.T, self.N
        # ppo_atari_oc.py:451-459 (same shapes and dtypes); one extra row so that row T receives the observation that
        # becomes `next_obs` (the bootstrap input of :534 and row 0 of the next iteration)
        self._obs = torch.zeros((T + 1, N, F, D), dtype=torch.float32, device=dev)
        self._dones = torch.zeros((T + 1, N), dtype=torch.float32, device=dev)
        self.actions = torch.zeros((T, N), dtype=torch.long, device=dev)
        self.logprobs = torch.zeros((T, N), dtype=torch.float32, device=dev)
        self.rewards = torch.zeros((T, N), dtype=torch.float32, device=dev)
        self.values = torch.zeros((T, N), dtype=torch.float32, device=dev)
        self._entropy = torch.empty(N, dtype=torch.float32, device=dev)          # computed by the kernel, unused (:506)
        # persistent pinned staging (the reference allocates pageable tensors every step)
        self._h_obs = torch.empty((N, F, D), dtype=torch.float32, pin_memory=True)
        self._h_reward = torch.empty(N, dtype=torch.float32, pin_memory=True)
        self._h_done = torch.empty(N, dtype=torch.float32, pin_memory=True)
        self._h_action = torch.empty(N, dtype=torch.int64, pin_memory=True)
        self.use_graphs = bool(use_graphs)
        self.feed = feed
        if feed is not None and tuple(feed.obs_shape) != (self.N, F, D):
            raise _abi.OcppoError(f"feed produces {tuple(feed.obs_shape)}, the agent expects {(self.N, F, D)}")
        self._carry_pending = False               # rows T hold next_obs / next_done that row 0 has not received yet
        self._action_ready = torch.cuda.Event()
        self._staging_free = torch.cuda.Event()
        self._staging_free.record(torch.cuda.current_stream(dev))
        agent._ensure_workspace(N, training=False)

    # -- views with the script's names and shapes ---------------------------------
    @property
    def obs(self):
        return self._obs[:self.T]

    @property
    def dones(self):
        return self._dones[:self.T]

    @property
    def next_obs(self):
        return self._obs[self.T]

    @property
    def next_done(self):
        return self._dones[self.T]

    # -- ppo_atari_oc.py:463-465 -----------------------------------------------------
    def reset(self, first_obs):
        """First observation of the run (``envs.reset()``); ``next_done`` starts at zero."""
        self._upload_obs(first_obs, 0)
        self._dones[0].zero_()
        self._carry_pending = False

    def _upload_obs(self, obs_np, row):
        self._staging_free.synchronize()          # the previous upload has left the staging buffers
        src = np.asarray(obs_np)
        if src.size != self._h_obs.numel() or src.shape[0] != self.N:
            raise _abi.OcppoError(f"observation batch {src.shape} does not match {tuple(self._h_obs.shape)}")
        self._h_obs.numpy()[...] = src.reshape(self._h_obs.shape)   # host cast to fp32 like torch.tensor(..., dtype=float32) (:513)
        self._obs[row].copy_(self._h_obs, non_blocking=True)
        self._staging_free.record(torch.cuda.current_stream(self.device))

    # -- the same two entry points fed with OBJECT SLOTS ---------------------
    def _require_feed(self):
        if self.feed is None:
            raise _abi.OcppoError("this collector was built without a feed (pass feed=ObjectFeed(...))")
        return self.feed

    def reset_objects(self, xy, visible):
        """``envs.reset()`` in object slots: xy ``(N, slots, 2)``, visible ``(N, slots)``; every env's frame stack is
        filled with its initial state (ocatari/core.py:241-244)."""
        feed = self._require_feed()
        self._staging_free.synchronize()
        feed.push(xy, visible, refill=np.ones(self.N, dtype=np.uint8), out=self._h_obs.numpy())
        self._obs[0].copy_(self._h_obs, non_blocking=True)
        self._staging_free.record(torch.cuda.current_stream(self.device))
        self._dones[0].zero_()
        self._carry_pending = False

    def observe_objects(self, step: int, xy, visible, reward, done):
        """``envs.step`` results of `step` as the env workers read them off the emulator: object slots, raw float64
        rewards, episode-end flags.  The frame-stacked observation block is assembled by ``ocppo_feed_push`` directly in
        the pinned staging buffer (an env that finished its episode was auto-reset by the vector env, so its stack is
        refilled), the rewards pass through ``VecNormalize``'s running-return scaling (``ocppo_feed_normalize_reward``),
        and the three uploads of ppo_atari_oc.py:512-514 follow."""
        feed = self._require_feed()
        self._staging_free.synchronize()
        done = np.asarray(done).astype(np.uint8).reshape(-1)
        feed.push(xy, visible, refill=done, out=self._h_obs.numpy())
        feed.normalize_reward(reward, done, out_reward=self._h_reward.numpy(), out_done=self._h_done.numpy())
        self._obs[step + 1].copy_(self._h_obs, non_blocking=True)
        self.rewards[step].copy_(self._h_reward, non_blocking=True)
        self._dones[step + 1].copy_(self._h_done, non_blocking=True)
        self._staging_free.record(torch.cuda.current_stream(self.device))
        if step == self.T - 1:
            self._carry_pending = True

    # -- ppo_atari_oc.py:502-509, 511 -------------------------------------------------
    def act(self, step: int) -> np.ndarray:
        """Sample the actions of `step` from ``obs[step]``; logprob / value / action land in row `step` of the
        buffers.  Returns the actions as a host int64 array (a view of the pinned staging buffer, valid until the
        next call)."""
        if not 0 <= step < self.T:
            raise _abi.OcppoError(f"step {step} outside [0, {self.T})")
        if step == 0 and self._carry_pending:
            self.start_next_iteration()
        stream = torch.cuda.current_stream(self.device)
        # consumes torch's CUDA generator like Categorical.sample(); writes rows `step` of the buffers in place
        self.agent.policy_step_into(self._obs[step], self.actions[step], self.logprobs[step], self._entropy,
                                    self.values[step], graph=self.use_graphs)
        self._h_action.copy_(self.actions[step], non_blocking=True)
        self._action_ready.record(stream)
        self._action_ready.synchronize()          # the environment needs the actions now (:511)
        return self._h_action.numpy()

    # -- ppo_atari_oc.py:512-514 ------------------------------------------------------
    def observe(self, step: int, next_obs, reward, next_done):
        """Results of ``envs.step`` for `step`: reward into ``rewards[step]``; the new observation / done flags into
        row ``step + 1`` (they are ``obs[step + 1]`` / ``dones[step + 1]`` of :502-503, or ``next_obs`` /
        ``next_done`` after the last step)."""
        self._upload_obs(next_obs, step + 1)
        self._h_reward.numpy()[...] = np.asarray(reward, dtype=np.float32).reshape(-1)
        self._h_done.numpy()[...] = np.asarray(next_done, dtype=np.float32).reshape(-1)
        self.rewards[step].copy_(self._h_reward, non_blocking=True)
        self._dones[step + 1].copy_(self._h_done, non_blocking=True)
        self._staging_free.record(torch.cuda.current_stream(self.device))
        if step == self.T - 1:
            self._carry_pending = True

    def start_next_iteration(self):
        """Carry ``next_obs`` / ``next_done`` into row 0 (the reference does this implicitly: its loop variables
        survive the iteration, :502-503).  ``act(0)`` calls this by itself when an iteration has completed; calling it
        explicitly (after the learner has consumed ``next_obs``) is harmless."""
        if not self._carry_pending:
            return
        self._obs[0].copy_(self._obs[self.T], non_blocking=True)
        self._dones[0].copy_(self._dones[self.T], non_blocking=True)
        self._carry_pending = False
