"""Fused PPO learner: GAE + the 16-minibatch update of one iteration.

Mirrors the learner section of the reference script as two calls:

  * ``compute_gae``  -- ``cleanrl/ppo_atari_oc.py:533-547``
  * ``update``       -- ``cleanrl/ppo_atari_oc.py:549-617`` (flatten, per-epoch
    ``np.random.shuffle``, minibatch loss, backward, ``clip_grad_norm_``, Adam)
    and, with ``world_size > 1``, the flat-gradient all-reduce of
    ``cleanrl/ppo_atari_multigpu.py:360-374`` on the gradient arena.

Hyper-parameter names are the script's CLI names (``Args``, :63-190).  All
arithmetic is done by libocppo kernels through the C ABI; torch provides device
memory, streams and ``torch.distributed`` only.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _abi
from .architectures.ppo import PPObj


@dataclass
class LearnerArgs:
    """Subset of the script's ``Args`` the learner reads (same names/defaults)."""
    num_envs: int = 10
    num_steps: int = 128
    gamma: float = 0.99
    gae_lambda: float = 0.95
    num_minibatches: int = 4
    update_epochs: int = 4
    norm_adv: bool = True
    clip_coef: float = 0.1
    clip_vloss: bool = True
    ent_coef: float = 0.01
    vf_coef: float = 0.5
    max_grad_norm: float = 0.5
    target_kl: Optional[float] = None
    learning_rate: float = 2.5e-4
    anneal_lr: bool = True
    # Adam as constructed at ppo_atari_oc.py:444
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_eps: float = 1e-5

    @property
    def batch_size(self):
        return int(self.num_envs * self.num_steps)

    @property
    def minibatch_size(self):
        return int(self.batch_size // self.num_minibatches)


def compute_gae(rewards, values, dones, next_value, next_done, gamma, gae_lambda, out=None, overlap_previous=False):
    """Stateless GAE on (T,N) fp32 CUDA tensors -> (advantages, returns).  ``overlap_previous``: launch the scan as a
    programmatic dependent of the kernel before it on the stream (``ocppo_gae_overlapped``) -- only when that kernel
    writes none of the five buffers but ``next_value``."""
    for t in (rewards, values, dones, next_value, next_done):
        if not t.is_cuda or t.dtype != torch.float32:
            raise _abi.OcppoError("compute_gae needs fp32 CUDA tensors (no CPU path)")
    T, N = rewards.shape
    rewards, values, dones = rewards.contiguous(), values.contiguous(), dones.contiguous()
    next_value = next_value.reshape(-1).contiguous()
    next_done = next_done.reshape(-1).contiguous()
    if next_value.numel() != N or next_done.numel() != N:
        raise _abi.OcppoError("next_value / next_done must have num_envs elements")
    if out is None:
        advantages, returns = torch.empty_like(rewards), torch.empty_like(rewards)
    else:
        advantages, returns = out
    stream = C.c_void_p(torch.cuda.current_stream(rewards.device).cuda_stream)
    fn = _abi.lib().ocppo_gae_overlapped if overlap_previous else _abi.lib().ocppo_gae
    _abi.check(fn(_abi.ptr(rewards), _abi.ptr(values), _abi.ptr(dones), _abi.ptr(next_value),
          _abi.ptr(next_done), T, N, float(gamma), float(gae_lambda), _abi.ptr(advantages), _abi.ptr(returns), stream))
    return advantages, returns


def make_permutations(batch_size, update_epochs, rng=None, pinned=True):
    """The reference's minibatch order: ``b_inds = np.arange(batch_size)`` then
    one cumulative ``np.random.shuffle`` per epoch (ppo_atari_oc.py:558-561).
    Returns an (update_epochs, batch_size) int64 host tensor (pinned)."""
    shuffle = np.random.shuffle if rng is None else rng.shuffle
    out = torch.empty((update_epochs, batch_size), dtype=torch.int64,
                      pin_memory=bool(pinned and torch.cuda.is_available()))
    view = out.numpy()
    b_inds = np.arange(batch_size)
    for epoch in range(update_epochs):
        shuffle(b_inds)
        view[epoch] = b_inds
    return out


class PPOLearner:
    """``use_graphs``: the ~45 launches of one minibatch (gather, forward, loss head, backward, reductions, clip + Adam)
    are captured into a CUDA graph the first time a minibatch slot is seen with a given set of buffer pointers and
    replayed afterwards -- the rollout buffers, the permutation buffer and the statistics rows persist across
    iterations, so after the first update every minibatch is one graph launch.  What changes from step to step (learning
    rate, Adam bias corrections) lives in device memory (``ocppo_adam_scalars_set`` / ``ocppo_clip_adam_dev``).  With
    ``world_size > 1`` the gradient all-reduce sits between two graphs (``graph_collective=False``, the default) or is
    captured with them.  If the caller hands over new buffers every iteration the graphs never hit and the learner
    falls back to plain launches by itself."""

    def __init__(self, agent: PPObj, args: LearnerArgs, world_size: int = 1, process_group=None,
                 chunk_rows: int = 32768, np_rng=None, use_graphs: bool = True, graph_collective: bool = False):
        if agent._handle is None:
            raise _abi.OcppoError("move the agent to its CUDA device before building the learner")
        self.agent = agent
        self.args = args
        self.world_size = int(world_size)
        self.process_group = process_group
        self.np_rng = np_rng
        self.device = agent._cuda_device
        self.adam_step = 0                      # torch Adam's `step`, never reset
        self.chunk_rows = int(chunk_rows)
        self.use_graphs = bool(use_graphs)
        self.graph_collective = bool(graph_collective)
        self._graphs = {}
        self._graph_hits = self._graph_misses = 0
        self._stats = self._gn = self._adam_dev = None
        self._perm_host = self._perm_dev = None
        self._gae_out = None
        self._ensure_slots(args.update_epochs * (args.num_minibatches + 1))
        self._hp = _abi.HParams(float(args.clip_coef), float(args.ent_coef), float(args.vf_coef),
                                int(bool(args.norm_adv)), int(bool(args.clip_vloss)))
        self.lr = float(args.learning_rate)
        self.launches = 0

    def _ensure_slots(self, n):
        if self._stats is None or self._stats.shape[0] < n:
            self._stats = torch.zeros((n, _abi.OCPPO_NUM_STATS), dtype=torch.float32, device=self.device)
            self._gn = torch.zeros(n, dtype=torch.float32, device=self.device)
            self._adam_dev = torch.zeros((n, 8), dtype=torch.float32, device=self.device)
            self._graphs.clear()

    # -- ppo_atari_oc.py:471-474 -------------------------------------------------
    def anneal_lr(self, iteration: int, num_iterations: int):
        if self.args.anneal_lr:
            frac = 1.0 - (iteration - 1.0) / num_iterations
            self.lr = frac * self.args.learning_rate
        return self.lr

    # -- ppo_atari_oc.py:533-547 -------------------------------------------------
    def compute_gae(self, rewards, values, dones, next_obs, next_done):
        """-> (advantages, returns), two learner-owned (T, N) buffers that the next call overwrites (the script
        allocates fresh ones every iteration, :535,:547; persistent buffers keep the update's CUDA graphs valid)."""
        with torch.no_grad():
            next_value = self.agent.get_value(next_obs).reshape(-1)
        if self._gae_out is None or self._gae_out[0].shape != rewards.shape:
            self._gae_out = (torch.empty_like(rewards), torch.empty_like(rewards))
        # the kernel before the scan is get_value's head kernel, whose only output is next_value: the scan may overlap it
        return compute_gae(rewards, values, dones, next_value, next_done, self.args.gamma, self.args.gae_lambda,
                           out=self._gae_out, overlap_previous=True)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- ppo_atari_oc.py:620-623 -------------------------------------------------
    def explained_variance(self, values, returns):
        """``1 - var(returns - values) / var(returns)`` over the flat rollout (nan when ``var(returns) == 0``), with
        both variances reduced on the device: 8 bytes come back instead of the two (T, N) buffers."""
        v, r = values.reshape(-1).contiguous(), returns.reshape(-1).contiguous()
        if v.numel() != r.numel() or not v.is_cuda or v.dtype != torch.float32 or r.dtype != torch.float32:
            raise _abi.OcppoError("explained_variance needs two fp32 CUDA tensors of equal size")
        self.agent._ensure_workspace(1, training=True)
        out = torch.empty(2, dtype=torch.float32, device=self.device)
        _abi.check(_abi.lib().ocppo_explained_variance(self.agent._handle.raw, _abi.ptr(v), _abi.ptr(r), v.numel(),
                                                       _abi.ptr(out), self._stream()))
        var_y, var_d = (float(x) for x in out.cpu())
        return float("nan") if var_y == 0 else 1.0 - var_d / var_y

    def minibatch_step(self, flat, idx_dev, slot):
        """One minibatch (ppo_atari_oc.py:563-610) on device index tensor `idx_dev`."""
        b_obs, b_actions, b_logprobs, b_advantages, b_returns, b_values = flat
        a = self.args
        mb = idx_dev.numel()
        self.agent._ensure_workspace(min(mb, self.chunk_rows), training=True, exact=True)
        lib, h = _abi.lib(), self.agent._handle.raw
        self.adam_step += 1
        stats, gn = self._stats[slot], self._gn[slot:slot + 1]

        def fwd_bwd(stream):
            _abi.check(lib.ocppo_minibatch_fwd_bwd(h, _abi.ptr(b_obs), _abi.ptr(b_actions), _abi.ptr(b_logprobs),
                                                   _abi.ptr(b_advantages), _abi.ptr(b_returns), _abi.ptr(b_values),
                                                   _abi.ptr(idx_dev), mb, C.byref(self._hp), _abi.ptr(stats), stream))

        def allreduce():
            # ppo_atari_multigpu.py:360-367: one SUM all-reduce of the flat gradient
            torch.distributed.all_reduce(self.agent.grad_arena, op=torch.distributed.ReduceOp.SUM,
                                         group=self.process_group)

        if not self.use_graphs:
            stream = self._stream()
            fwd_bwd(stream)
            if self.world_size > 1:
                allreduce()
            _abi.check(lib.ocppo_clip_adam(h, 1.0 / self.world_size, float(a.max_grad_norm), float(self.lr),
                                           a.adam_beta1, a.adam_beta2, a.adam_eps, self.adam_step, _abi.ptr(gn), stream))
            return

        scal = self._adam_dev[slot]
        _abi.check(lib.ocppo_adam_scalars_set(_abi.ptr(scal), 1.0 / self.world_size, float(a.max_grad_norm),
                                              float(self.lr), a.adam_beta1, a.adam_beta2, a.adam_eps, self.adam_step,
                                              self._stream()))

        def adam(stream):
            _abi.check(lib.ocppo_clip_adam_dev(h, _abi.ptr(scal), _abi.ptr(gn), stream))

        key = (tuple(t.data_ptr() for t in flat), idx_dev.data_ptr(), mb, slot, self.agent._workspace.data_ptr(),
               self.agent._ws_rows, lib.ocppo_get_gemm_mode())
        entry = self._graphs.get(key)
        if entry is None:
            self._graph_misses += 1
            if self._graph_misses > 256 and self._graph_hits < self._graph_misses:
                self.use_graphs = False      # the caller's buffers never repeat: capturing costs more than it saves
                self._graphs.clear()
            elif len(self._graphs) >= 256:
                self._graphs.clear()
            # the captured sequence must not depend on whether a rollout left prepared weights behind
            _abi.check(lib.ocppo_weights_version(h, -1))
            dev = self.device
            if self.world_size > 1 and not self.graph_collective:
                entry = (_abi.capture(h, dev, fwd_bwd), _abi.capture(h, dev, adam))
            else:
                def whole(stream):
                    fwd_bwd(stream)
                    if self.world_size > 1:
                        allreduce()
                    adam(stream)
                entry = (_abi.capture(h, dev, whole),)
            self._graphs[key] = entry
        else:
            self._graph_hits += 1
        entry[0].replay()
        if len(entry) == 2:
            allreduce()
            entry[1].replay()

    def update(self, obs, actions, logprobs, advantages, returns, values, permutations=None):
        """ppo_atari_oc.py:549-617.  Buffers are the (T, N, ...) rollout tensors.
        Returns a dict of python floats like the script's logging (:612-614,:624-641)."""
        a = self.args
        batch_size = logprobs.numel()
        minibatch_size = int(batch_size // a.num_minibatches)
        flat = (obs.reshape(batch_size, -1).contiguous(), actions.reshape(-1).contiguous(),
                logprobs.reshape(-1).contiguous(), advantages.reshape(-1).contiguous(),
                returns.reshape(-1).contiguous(), values.reshape(-1).contiguous())
        if flat[1].dtype != torch.int64:
            raise _abi.OcppoError("actions must be int64 (torch.long), as in the reference buffers")
        # Without given permutations the reference's order is produced here, one cumulative host shuffle per epoch
        # (:558-561), each into its own pinned row and copied asynchronously: the shuffle of epoch e+1 runs on the
        # host while the device is busy with epoch e.
        lazy = permutations is None
        if self._perm_dev is None or tuple(self._perm_dev.shape) != (a.update_epochs, batch_size):
            self._perm_host = torch.empty((a.update_epochs, batch_size), dtype=torch.int64, pin_memory=True)
            self._perm_dev = torch.empty((a.update_epochs, batch_size), dtype=torch.int64, device=self.device)
        perm_dev = self._perm_dev            # persistent: the minibatch graphs are keyed on its rows
        if lazy:
            shuffle = np.random.shuffle if self.np_rng is None else self.np_rng.shuffle
            b_inds = np.arange(batch_size)
        else:
            perm_dev.copy_(permutations, non_blocking=True)
        self._ensure_slots(a.update_epochs * (-(-batch_size // minibatch_size)))
        slot, epochs_run = 0, 0
        for epoch in range(a.update_epochs):
            if lazy:
                shuffle(b_inds)
                self._perm_host[epoch].numpy()[:] = b_inds
                perm_dev[epoch].copy_(self._perm_host[epoch], non_blocking=True)
            for start in range(0, batch_size, minibatch_size):
                end = start + minibatch_size
                idx = perm_dev[epoch, start:end]
                if idx.numel() == 0:
                    continue
                self.minibatch_step(flat, idx, slot)
                slot += 1
            epochs_run += 1
            if a.target_kl is not None:
                approx_kl = float(self._stats[slot - 1, 5].item())
                if approx_kl > a.target_kl:
                    break
        stats = self._stats[:slot].cpu()
        gn = self._gn[:slot].cpu()
        last = stats[slot - 1]
        out = {name: float(last[i]) for i, name in enumerate(_abi.STAT_NAMES)}
        out["clipfracs"] = [float(x) for x in stats[:, 6]]
        out["grad_norm"] = float(gn[slot - 1])
        out["grad_norms"] = [float(x) for x in gn]
        out["minibatches"] = slot
        out["epochs"] = epochs_run
        return out
