"""Drop-in ``PPObj`` (reference: ``cleanrl/architectures/ppo.py:60-95``).

Same constructor, ``get_value`` / ``get_action_and_value`` surface, parameter
names and ``parameters()`` order as the reference class, so
``cleanrl/ppo_atari_oc.py:438-444`` (``Agent(envs, device, encoder_dims,
decoder_dims).to(device)`` + ``optim.Adam(agent.parameters())``) works unchanged.

Differences in mechanism, not in behaviour:
  * once the module sits on a CUDA device its parameters become views of one
    flat fp32 arena (``self.arena``), the buffer the CUDA kernels read and the
    buffer a data-parallel learner all-reduces (ppo_atari_multigpu.py:360-374);
  * forward passes run in ``libocppo.so``; under ``torch.no_grad`` the whole
    of ppo.py:89-95 (MLP, logits, Categorical sample, log-prob, entropy, value)
    is the fused rollout path, and sampling consumes torch's CUDA Philox
    generator exactly like ``Categorical.sample()`` would;
  * with grad enabled the trunk is a ``torch.autograd.Function`` over the same
    kernels (compat tier, so the script's own loss/backward/Adam keep working).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

import torch
import torch.nn as nn
from torch.distributions.categorical import Categorical

from .. import _abi
from .common import Predictor, layer_init  # noqa: F401  (re-export like the reference)


class _TrunkFunction(torch.autograd.Function):
    """hidden = network(x) through libocppo, differentiable wrt the trunk params."""

    @staticmethod
    def forward(ctx, agent, x, *params):
        B = x.shape[0]
        agent._ensure_workspace(B, training=True)
        agent._hint_weights()
        hidden = torch.empty((B, agent._hidden), dtype=torch.float32, device=x.device)
        _abi.check(_abi.lib().ocppo_trunk_forward(agent._handle.raw, _abi.ptr(x), B, _abi.ptr(hidden),
                                                 agent._stream()))
        ctx.agent = agent
        ctx.save_for_backward(x)
        ctx.version = agent._fwd_version = agent._fwd_version + 1
        return hidden

    @staticmethod
    def backward(ctx, d_hidden):
        agent = ctx.agent
        if ctx.version != agent._fwd_version:
            raise RuntimeError("PPObj compat tier keeps the activations of the latest grad-mode forward only")
        (x,) = ctx.saved_tensors
        d_hidden = d_hidden.contiguous()
        _abi.check(_abi.lib().ocppo_trunk_backward(agent._handle.raw, _abi.ptr(x), x.shape[0],
                                                  _abi.ptr(d_hidden), agent._stream()))
        grads = [agent._grad_view(i).clone() for i in range(agent._n_trunk_tensors)]
        return (None, None, *grads)


class PPObj(Predictor):
    def __init__(self, envs, device, encoder_dims=(128, 64), decoder_dims=(32,)):
        super().__init__()
        self.device = device
        dims = envs.observation_space.shape
        self._frames = int(np.prod(dims[:-1], dtype=int))
        self._obs_dim = int(dims[-1])
        self._n_actions = int(envs.action_space.n)
        self._encoder_dims = tuple(int(d) for d in encoder_dims)
        self._decoder_dims = tuple(int(d) for d in decoder_dims)

        # Same module tree and RNG consumption as the reference (ppo.py:66-84).
        mods = []
        width = self._obs_dim
        for out in self._encoder_dims:
            mods += [layer_init(nn.Linear(width, out)), nn.ReLU()]
            width = out
        mods.append(nn.Flatten())
        width *= self._frames
        for out in self._decoder_dims:
            mods += [layer_init(nn.Linear(width, out)), nn.ReLU()]
            width = out
        self.network = nn.Sequential(*mods)
        self.actor = layer_init(nn.Linear(width, self._n_actions), std=0.01)
        self.critic = layer_init(nn.Linear(width, 1), std=1)
        self._hidden = width
        self._n_trunk_tensors = 2 * (len(self._encoder_dims) + len(self._decoder_dims))

        self._handle = None
        self.arena = None           # flat fp32 parameters (padded), device
        self.grad_arena = None
        self.exp_avg = None
        self.exp_avg_sq = None
        self._workspace = None
        self._ws_rows = 0
        self._ws_training = False
        self._fwd_version = 0
        self._manual_version = 0
        # forward passes over more rows than this run in chunks inside the library (ocppo_policy_step / ocppo_value), so
        # one large no-grad call never sizes the workspace -- and with it the training chunk -- by its batch
        self.inference_chunk_rows = 65536
        self._graphs = {}           # CUDA graphs of the fused rollout step, keyed by the pointers they were captured on
        self._philox_dev = None     # (seed, offset) the graph-replayed head kernel reads

    # ---- arena management ---------------------------------------------------
    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        p = next(self.parameters())
        if p.is_cuda:
            self._materialize(p.device)
        return out

    def _materialize(self, dev):
        """Move the parameters into one flat device arena and re-point them."""
        for p in self.parameters():
            if p.dtype != torch.float32:
                raise _abi.OcppoError("oc_cleanrl_b200.PPObj computes in fp32 only")
        cfg = _abi.make_config(self._frames, self._obs_dim, self._n_actions, self._encoder_dims,
                               self._decoder_dims, dev.index if dev.index is not None else torch.cuda.current_device())
        self._handle = _abi.Handle(cfg)
        h = self._handle
        params = list(self.parameters())
        if len(params) != h.n_tensors:
            raise _abi.OcppoError("parameter list does not match the C layout")
        with torch.no_grad():
            arena = torch.zeros(h.arena_floats, dtype=torch.float32, device=dev)
            for p, (off, rows, cols) in zip(params, h.layout):
                if p.numel() != rows * cols:
                    raise _abi.OcppoError("parameter shape does not match the C layout")
                view = arena[off:off + rows * cols].view(p.shape)
                view.copy_(p.data)
                p.data = view
        self.arena = arena
        self.grad_arena = torch.zeros_like(arena)
        self.exp_avg = torch.zeros_like(arena)
        self.exp_avg_sq = torch.zeros_like(arena)
        _abi.check(_abi.lib().ocppo_bind_arena(h.raw, _abi.ptr(self.arena), _abi.ptr(self.grad_arena),
                                               _abi.ptr(self.exp_avg), _abi.ptr(self.exp_avg_sq)))
        self._workspace, self._ws_rows, self._ws_training = None, 0, False
        props = torch.cuda.get_device_properties(dev)
        self._sm_count = props.multi_processor_count
        self._max_threads_per_sm = props.max_threads_per_multi_processor
        self._cuda_device = dev

    def _require_handle(self):
        if self._handle is None:
            raise _abi.OcppoError(
                "oc_cleanrl_b200.PPObj has no CPU path: move the agent to a CUDA (sm_100) device first")
        return self._handle

    def _ensure_workspace(self, rows, training, exact=False):
        """Bind a workspace for chunks of `rows` samples.

        * inference (``training=False``): the library chunks ``ocppo_policy_step`` / ``ocppo_value`` by whatever is bound,
          so a bound TRAINING workspace is used as it is, and otherwise the request is capped at ``inference_chunk_rows``
          and only grows the binding -- one large no-grad call never sizes the workspace by its batch;
        * ``exact=True`` (the fused learner): binds exactly `rows`, so the update runs with the chunk that was asked for,
          whatever forward calls came before;
        * training without ``exact`` (the autograd bridge): grows to at least `rows`."""
        h = self._require_handle()
        rows = int(rows)
        have = self._workspace is not None
        if not training:
            if have and self._ws_training:
                return
            rows = min(rows, int(self.inference_chunk_rows))
            if have and rows <= self._ws_rows:
                return
            rows = max(rows, self._ws_rows)
        elif exact:
            if have and self._ws_training and self._ws_rows == rows:
                return
        else:
            if have and self._ws_training and rows <= self._ws_rows:
                return
            rows = max(rows, self._ws_rows)
        nbytes = h.workspace_bytes(rows, training)
        self._workspace = None
        self._graphs.clear()        # captured launches point into the old workspace
        ws = torch.empty(nbytes, dtype=torch.uint8, device=self._cuda_device)
        _abi.check(_abi.lib().ocppo_bind_workspace(h.raw, _abi.ptr(ws), nbytes, rows, int(training)))
        self._workspace, self._ws_rows, self._ws_training = ws, rows, bool(training)

    @property
    def workspace_rows(self):
        """Rows per forward / backward chunk of the bound workspace (0: none bound yet)."""
        return self._ws_rows

    @property
    def workspace_bytes(self):
        return 0 if self._workspace is None else int(self._workspace.numel())

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self._cuda_device).cuda_stream)

    def _grad_view(self, i):
        off, rows, cols = self._handle.layout[i]
        p = list(self.parameters())[i]
        return self.grad_arena[off:off + rows * cols].view(p.shape)

    def _check_obs(self, x):
        if not x.is_cuda:
            raise _abi.OcppoError("observations must be on the agent's CUDA device (no CPU path)")
        if x.dtype != torch.float32:
            x = x.float()
        x = x.contiguous()
        if x.dim() < 2 or int(np.prod(x.shape[1:])) != self._frames * self._obs_dim:
            raise _abi.OcppoError(f"expected observations (B,{self._frames},{self._obs_dim}), got {tuple(x.shape)}")
        return x

    def _philox_state(self, numel):
        """Consume torch's CUDA generator the way ``exponential_`` of `numel`
        elements would (ATen/native/cuda/DistributionTemplates.h:50-62)."""
        gen = torch.cuda.default_generators[self._cuda_device.index]
        seed, offset = gen.initial_seed(), gen.get_offset()
        inc = _abi.lib().ocppo_philox_increment(numel, self._sm_count, self._max_threads_per_sm)
        gen.set_offset(offset + inc)
        return seed, offset

    # ---- public surface (architectures/ppo.py:86-95) ---------------------------
    def _needs_grad(self):
        return torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())

    def _hidden_with_grad(self, x):
        trunk = list(self.parameters())[: self._n_trunk_tensors]
        return _TrunkFunction.apply(self, x, *trunk)

    def _hint_weights(self):
        """Tell the library whether the parameter arena changed since the last call (``ocppo_weights_version``): the
        rollout loop (ppo_atari_oc.py:500-514) calls the policy 128 times on unchanged weights, and the FP16 engines
        then skip their per-call weight preparation.  The counter is the sum of torch's version counters of the parameters,
        which every in-place op on a parameter (optimizer steps, ``load_state_dict``, ``copy_``) increments; writes
        through ``param.data`` or straight into ``self.arena`` bypass it -- call :meth:`weights_changed` after those."""
        # the parameters were re-pointed at the arena through ``.data`` (which keeps each parameter's own version
        # counter), so the arena's change counter is the sum of theirs
        version = sum(p._version for p in self.parameters()) + self._manual_version
        _abi.check(_abi.lib().ocppo_weights_version(self._handle.raw, int(version)))

    def weights_changed(self):
        """Invalidate the prepared weights by hand (only needed after writes torch does not track, e.g. ``p.data``)."""
        self._manual_version += 1 << 40

    def get_value(self, x):
        x = self._check_obs(x)
        if self._needs_grad():
            return self.critic(self._hidden_with_grad(x))
        B = x.shape[0]
        self._ensure_workspace(B, training=False)
        self._hint_weights()
        value = torch.empty((B, 1), dtype=torch.float32, device=x.device)
        _abi.check(_abi.lib().ocppo_value(self._handle.raw, _abi.ptr(x), B, _abi.ptr(value), self._stream()))
        return value

    def get_action_and_value(self, x, action=None):
        x = self._check_obs(x)
        if self._needs_grad():
            hidden = self._hidden_with_grad(x)
            logits = self.actor(hidden)
            probs = Categorical(logits=logits)
            if action is None:
                action = probs.sample()
            return action, probs.log_prob(action), probs.entropy(), self.critic(hidden)
        B = x.shape[0]
        dev = x.device
        out_action = torch.empty(B, dtype=torch.int64, device=dev)
        logp = torch.empty(B, dtype=torch.float32, device=dev)
        entropy = torch.empty(B, dtype=torch.float32, device=dev)
        value = torch.empty((B, 1), dtype=torch.float32, device=dev)
        self.policy_step_into(x, out_action, logp, entropy, value, given=action)
        return out_action, logp, entropy, value

    def policy_step_into(self, x, action, logp, entropy, value, given=None, graph=False):
        """ppo.py:89-95 under ``no_grad`` with caller-owned outputs (int64 (B,), fp32 (B,), fp32 (B,), fp32 (B,) or
        (B, 1)): the fused kernel writes e.g. ``actions[step]`` / ``logprobs[step]`` / ``values[step]`` of the rollout
        buffers in place (ppo_atari_oc.py:506-509 are three copies).  ``graph=True`` replays the launches of the call from
        a CUDA graph captured on first use for these pointers (buffers that persist across iterations, as the rollout
        buffers do); the Philox (seed, offset) pair then lives in device memory and is refreshed before each replay."""
        B = x.shape[0]
        self._ensure_workspace(B, training=False)
        self._hint_weights()
        lib, h = _abi.lib(), self._handle.raw
        seed = offset = 0
        if given is None:
            seed, offset = self._philox_state(B * self._n_actions)
        else:
            given = given.to(device=x.device, dtype=torch.int64).contiguous()
        if not graph or given is not None:
            _abi.check(lib.ocppo_policy_step(h, _abi.ptr(x), B, seed, offset, _abi.ptr(given), _abi.ptr(action),
                                             _abi.ptr(logp), _abi.ptr(entropy), _abi.ptr(value), self._stream()))
            return
        if self._philox_dev is None:
            self._philox_dev = torch.zeros(2, dtype=torch.int64, device=self._cuda_device)
        _abi.check(lib.ocppo_philox_set(_abi.ptr(self._philox_dev), seed, offset, self._stream()))
        # the launch sequence depends on whether the prepared weights can be reused (ocppo_weights_version), so that
        # state is part of the key: a graph captured with the preparation in it is replayed only when it is needed
        prepared = bool(lib.ocppo_weights_prepared(h))
        key = ("policy", x.data_ptr(), B, action.data_ptr(), logp.data_ptr(), entropy.data_ptr(), value.data_ptr(), prepared,
               lib.ocppo_get_gemm_mode())
        g = self._graphs.get(key)
        if g is None:
            if len(self._graphs) >= 1024:
                self._graphs.clear()
            g = self._graphs[key] = _abi.capture(h, self._cuda_device, lambda stream: _abi.check(
                lib.ocppo_policy_step_dev(h, _abi.ptr(x), B, _abi.ptr(self._philox_dev), None, _abi.ptr(action),
                                          _abi.ptr(logp), _abi.ptr(entropy), _abi.ptr(value), stream)))
        g.replay()
