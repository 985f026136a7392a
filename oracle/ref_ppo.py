"""CPU restatement (torch, fp32) of the reference's PPO_OBJ learner path.

TEST INFRASTRUCTURE -- see ``oracle/__init__.py`` for who may import this.

Every function cites the reference lines it follows (paths relative to
``/root/reference``).  The arithmetic is executed by the reference's own
dependency (torch eager ops, in the reference's operator order); nothing here
is re-derived math.
"""
from __future__ import annotations

import os
import sys
from dataclasses import dataclass
from types import SimpleNamespace
from typing import Optional, Sequence

import numpy as np
import torch
import torch.nn as nn
from torch.distributions.categorical import Categorical

REFERENCE_ROOT = "/root/reference"


# --------------------------------------------------------------------------
# hyper-parameters: cleanrl/ppo_atari_oc.py:63-190 (defaults of THAT script)
# --------------------------------------------------------------------------
@dataclass
class PPOArgs:
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
    seed: int = 42
    encoder_dims: Sequence[int] = (256, 512, 1024, 512)
    decoder_dims: Sequence[int] = (512,)

    # cleanrl/ppo_atari_oc.py:344-346
    @property
    def batch_size(self) -> int:
        return int(self.num_envs * self.num_steps)

    @property
    def minibatch_size(self) -> int:
        return int(self.batch_size // self.num_minibatches)


def stub_envs(frames: int, obs_dim: int, n_actions: int):
    """The three attributes PPObj.__init__ reads (architectures/ppo.py:65,83)."""
    return SimpleNamespace(
        observation_space=SimpleNamespace(shape=(frames, obs_dim)),
        action_space=SimpleNamespace(n=n_actions),
    )


# --------------------------------------------------------------------------
# model: cleanrl/architectures/ppo.py:9-13 (layer_init), :60-95 (PPObj),
#        cleanrl/architectures/common.py:13-16 (Predictor.predict)
# --------------------------------------------------------------------------
def _layer_init(layer, std=np.sqrt(2), bias_const=0.0):
    nn.init.orthogonal_(layer.weight, std)
    nn.init.constant_(layer.bias, bias_const)
    return layer


class OraclePPObj(nn.Module):
    """Restatement of ``PPObj``; consumes the torch RNG in the same order so
    that ``torch.manual_seed(s); OraclePPObj(...)`` reproduces the reference's
    initial weights bit for bit (checked in tests/test_oracle_cpu.py::test_oracle_model_equals_reference_model).
    """

    def __init__(self, envs, device="cpu", encoder_dims=(128, 64), decoder_dims=(32,)):
        super().__init__()
        self.device = device
        dims = envs.observation_space.shape
        mods = []
        width = dims[-1]
        for out in encoder_dims:                       # ppo.py:70-73
            mods += [_layer_init(nn.Linear(width, out)), nn.ReLU()]
            width = out
        mods.append(nn.Flatten())                      # ppo.py:74
        width *= int(np.prod(dims[:-1], dtype=int))    # ppo.py:75
        last = width
        for out in decoder_dims:                       # ppo.py:77-80
            mods += [_layer_init(nn.Linear(width, out)), nn.ReLU()]
            width = out
            last = out
        self.network = nn.Sequential(*mods)
        self.actor = _layer_init(nn.Linear(last, envs.action_space.n), std=0.01)   # ppo.py:83
        self.critic = _layer_init(nn.Linear(last, 1), std=1)                       # ppo.py:84

    def get_value(self, x):                            # ppo.py:86-87
        return self.critic(self.network(x))

    def get_action_and_value(self, x, action=None):    # ppo.py:89-95
        hidden = self.network(x)
        logits = self.actor(hidden)
        probs = Categorical(logits=logits)
        if action is None:
            action = probs.sample()
        return action, probs.log_prob(action), probs.entropy(), self.critic(hidden)

    def predict(self, x, states=None, **_):            # common.py:13-16
        with torch.no_grad():
            z = self.actor(self.network(torch.Tensor(x).to(self.device)))
            return np.argmax(z.cpu().numpy(), axis=1), states


def import_reference_ppobj():
    """The reference's own class; only possible where /root/reference exists."""
    path = os.path.join(REFERENCE_ROOT, "cleanrl")
    if not os.path.isdir(path):
        return None
    if path not in sys.path:
        sys.path.insert(0, path)
    from architectures.ppo import PPObj  # type: ignore
    return PPObj


def make_agent(frames, obs_dim, n_actions, encoder_dims, decoder_dims, seed=42,
               use_reference=False, dtype=torch.float32):
    """Seeded construction as cleanrl/ppo_atari_oc.py:359,438-440."""
    torch.manual_seed(seed)
    cls = import_reference_ppobj() if use_reference else OraclePPObj
    if cls is None:
        raise RuntimeError("reference tree not present")
    agent = cls(stub_envs(frames, obs_dim, n_actions), "cpu", tuple(encoder_dims), tuple(decoder_dims))
    return agent.to(dtype)


# --------------------------------------------------------------------------
# GAE: cleanrl/ppo_atari_oc.py:533-547 (== ppo_atari_multigpu.py:288-301)
# --------------------------------------------------------------------------
@torch.no_grad()
def ref_gae(rewards, values, dones, next_value, next_done, gamma, gae_lambda):
    """rewards/values/dones (T,N); next_value (1,N) or (N,); next_done (N,)."""
    num_steps = rewards.shape[0]
    next_value = next_value.reshape(1, -1)
    advantages = torch.zeros_like(rewards)
    lastgaelam = 0.0
    for t in reversed(range(num_steps)):
        if t == num_steps - 1:
            nextnonterminal = 1.0 - next_done
            nextvalues = next_value
        else:
            nextnonterminal = 1.0 - dones[t + 1]
            nextvalues = values[t + 1]
        delta = rewards[t] + gamma * nextvalues * nextnonterminal - values[t]
        lastgaelam = delta + gamma * gae_lambda * nextnonterminal * lastgaelam
        advantages[t] = lastgaelam
    returns = advantages + values
    return advantages, returns


def np_gae_fp32(rewards, values, dones, next_value, next_done, gamma, gae_lambda):
    """Explicit-association fp32 numpy restatement; the
    device kernel follows exactly this order, so equality with it is bitwise."""
    f = np.float32
    r, v, d = (np.asarray(a, dtype=f) for a in (rewards, values, dones))
    T, N = r.shape
    g, c = f(gamma), f(gamma * gae_lambda)
    adv = np.zeros((T, N), f)
    last = np.zeros(N, f)
    nv = np.asarray(next_value, f).reshape(-1)
    nnt = f(1.0) - np.asarray(next_done, f).reshape(-1)
    for t in range(T - 1, -1, -1):
        if t != T - 1:
            nv, nnt = v[t + 1], f(1.0) - d[t + 1]
        delta = (r[t] + ((g * nv) * nnt)) - v[t]
        last = delta + ((c * nnt) * last)
        adv[t] = last
    return adv, adv + v


# --------------------------------------------------------------------------
# loss head: cleanrl/ppo_atari_oc.py:566-602
# --------------------------------------------------------------------------
def ref_minibatch_loss(agent, b_obs, b_actions, b_logprobs, b_advantages, b_returns, b_values,
                       mb_inds, args: PPOArgs):
    """Returns (loss, stats dict of 0-d tensors)."""
    _, newlogprob, entropy, newvalue = agent.get_action_and_value(
        b_obs[mb_inds], b_actions.long()[mb_inds])
    logratio = newlogprob - b_logprobs[mb_inds]
    ratio = logratio.exp()

    with torch.no_grad():
        old_approx_kl = (-logratio).mean()
        approx_kl = ((ratio - 1) - logratio).mean()
        clipfrac = ((ratio - 1.0).abs() > args.clip_coef).float().mean()

    mb_advantages = b_advantages[mb_inds]
    if args.norm_adv:
        mb_advantages = (mb_advantages - mb_advantages.mean()) / (mb_advantages.std() + 1e-8)

    pg_loss1 = -mb_advantages * ratio
    pg_loss2 = -mb_advantages * torch.clamp(ratio, 1 - args.clip_coef, 1 + args.clip_coef)
    pg_loss = torch.max(pg_loss1, pg_loss2).mean()

    newvalue = newvalue.view(-1)
    if args.clip_vloss:
        v_loss_unclipped = (newvalue - b_returns[mb_inds]) ** 2
        v_clipped = b_values[mb_inds] + torch.clamp(
            newvalue - b_values[mb_inds], -args.clip_coef, args.clip_coef)
        v_loss_clipped = (v_clipped - b_returns[mb_inds]) ** 2
        v_loss_max = torch.max(v_loss_unclipped, v_loss_clipped)
        v_loss = 0.5 * v_loss_max.mean()
    else:
        v_loss = 0.5 * ((newvalue - b_returns[mb_inds]) ** 2).mean()

    entropy_loss = entropy.mean()
    loss = pg_loss - args.ent_coef * entropy_loss + v_loss * args.vf_coef
    stats = dict(pg_loss=pg_loss.detach(), v_loss=v_loss.detach(), entropy=entropy_loss.detach(),
                 loss=loss.detach(), old_approx_kl=old_approx_kl, approx_kl=approx_kl,
                 clipfrac=clipfrac)
    return loss, stats


def make_optimizer(agent, args: PPOArgs):
    """cleanrl/ppo_atari_oc.py:444."""
    return torch.optim.Adam(agent.parameters(), lr=args.learning_rate, eps=1e-5)


def anneal_lr(optimizer, iteration, num_iterations, args: PPOArgs):
    """cleanrl/ppo_atari_oc.py:471-474 (iteration is 1-based)."""
    if args.anneal_lr:
        frac = 1.0 - (iteration - 1.0) / num_iterations
        lrnow = frac * args.learning_rate
        optimizer.param_groups[0]["lr"] = lrnow
    return optimizer.param_groups[0]["lr"]


class relu_decisions:
    """Context manager over the trunk's ReLUs (test infrastructure).  Always records, per trunk layer, the
    Linear input ``x``, the pre-activation ``z`` and the ReLU decision ``z > 0``.  With ``masks`` (one bool
    tensor per trunk layer, broadcastable to ``z``) the ReLU output is replaced by ``z * mask``: the
    reference network evaluated with somebody else's decisions.  A decision is a step function of ``z``; two
    fp32 evaluations of the same dot product that differ in the last bit can decide a ``z`` at rounding
    distance from zero differently, and the backward pass then differs by that sample's whole contribution
    to the unit -- which no tolerance on the GEMM arithmetic can bound.  The GPU parity test therefore hands
    the reference the decisions of the path under test and checks separately that every decision that
    differs from the reference's own was within rounding of zero."""

    def __init__(self, agent, masks=None):
        self.agent, self.masks, self.records, self._hooks = agent, masks, [], []

    def __enter__(self):
        mods = list(self.agent.network)
        layer = 0
        for i, m in enumerate(mods):
            if isinstance(m, nn.Linear) and i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU):
                rec = {"linear": m}
                self.records.append(rec)
                self._hooks.append(m.register_forward_hook(
                    lambda mod, inp, out, rec=rec: rec.update(x=inp[0].detach(), z=out.detach())))
                mask = None if self.masks is None else self.masks[layer]
                def relu_hook(mod, inp, out, rec=rec, mask=mask):
                    rec["decision"] = inp[0].detach() > 0
                    if mask is not None:
                        return inp[0] * mask.reshape(inp[0].shape).to(inp[0].dtype)
                self._hooks.append(mods[i + 1].register_forward_hook(relu_hook))
                layer += 1
        return self

    def __exit__(self, *exc):
        for h in self._hooks:
            h.remove()
        return False


def flat_grads(agent):
    return torch.cat([p.grad.reshape(-1) for p in agent.parameters()])


def flat_params(agent):
    return torch.cat([p.detach().reshape(-1) for p in agent.parameters()])


def ref_minibatch_step(agent, optimizer, batch, mb_inds, args: PPOArgs, world_grads=None):
    """One minibatch: cleanrl/ppo_atari_oc.py:563-610.

    ``world_grads``: optional callable(flat_grad)->flat_grad applied between
    backward and clip (the allreduce hook of ppo_atari_multigpu.py:360-374).
    Returns (stats, grad_norm, flat_grad_before_clip).
    """
    loss, stats = ref_minibatch_loss(agent, *batch, mb_inds, args)
    for p in agent.parameters():          # :605-606
        p.grad = None
    loss.backward()                       # :608
    if world_grads is not None:
        g = world_grads(flat_grads(agent))
        off = 0
        for p in agent.parameters():
            p.grad.data.copy_(g[off:off + p.numel()].view_as(p.grad.data))
            off += p.numel()
    raw = flat_grads(agent).clone()
    gn = nn.utils.clip_grad_norm_(agent.parameters(), args.max_grad_norm)   # :609
    optimizer.step()                      # :610
    return stats, gn.detach(), raw


def ref_update(agent, optimizer, obs, actions, logprobs, advantages, returns, values,
               args: PPOArgs, np_rng=None, trace=None):
    """The update section, cleanrl/ppo_atari_oc.py:549-617.

    ``np_rng``: a ``np.random.RandomState`` standing for the script's global
    numpy RNG (``np.random.shuffle`` at :561 uses the legacy MT19937 stream).
    Returns (last-minibatch stats, clipfracs list, last grad norm).
    """
    obs_shape = obs.shape[2:]
    b_obs = obs.reshape((-1,) + tuple(obs_shape))          # :550
    b_logprobs = logprobs.reshape(-1)
    b_actions = actions.reshape(-1)
    b_advantages = advantages.reshape(-1)
    b_returns = returns.reshape(-1)
    b_values = values.reshape(-1)
    batch = (b_obs, b_actions, b_logprobs, b_advantages, b_returns, b_values)

    shuffle = np.random.shuffle if np_rng is None else np_rng.shuffle
    batch_size = b_logprobs.numel()
    minibatch_size = int(batch_size // args.num_minibatches)
    b_inds = np.arange(batch_size)                          # :558
    clipfracs, stats, gn = [], None, None
    for epoch in range(args.update_epochs):                 # :560
        shuffle(b_inds)                                     # :561
        for start in range(0, batch_size, minibatch_size):  # :562
            end = start + minibatch_size
            mb_inds = b_inds[start:end]
            stats, gn, raw = ref_minibatch_step(agent, optimizer, batch, mb_inds, args)
            clipfracs.append(stats["clipfrac"].item())
            if trace is not None:
                trace.append(dict(stats=stats, gn=gn, mb_inds=mb_inds.copy()))
        if args.target_kl is not None and stats["approx_kl"] > args.target_kl:   # :616-617
            break
    return stats, clipfracs, gn


def explained_variance(b_values, b_returns):
    """cleanrl/ppo_atari_oc.py:620-622."""
    y_pred, y_true = b_values.detach().cpu().numpy(), b_returns.detach().cpu().numpy()
    var_y = np.var(y_true)
    return np.nan if var_y == 0 else 1 - np.var(y_true - y_pred) / var_y


# --------------------------------------------------------------------------
# data-parallel variant: cleanrl/ppo_atari_multigpu.py:166-173,208-211,360-377
# --------------------------------------------------------------------------
def ref_dp_minibatch_step(agents, optimizers, batches, mb_inds_per_rank, args: PPOArgs):
    """In-process K-rank simulation of one DP minibatch: every rank computes
    its local gradient, gradients are summed, divided by world_size, then each
    rank clips and steps identically (ppo_atari_multigpu.py:360-377)."""
    world = len(agents)
    flats, stats_all = [], []
    for agent, batch, inds in zip(agents, batches, mb_inds_per_rank):
        loss, stats = ref_minibatch_loss(agent, *batch, inds, args)
        for p in agent.parameters():
            p.grad = None
        loss.backward()
        flats.append(flat_grads(agent))
        stats_all.append(stats)
    total = flats[0].clone()
    for g in flats[1:]:
        total = total + g                 # all_reduce(SUM), :367
    gns = []
    for agent, opt in zip(agents, optimizers):
        off = 0
        for p in agent.parameters():      # :368-374
            p.grad.data.copy_(total[off:off + p.numel()].view_as(p.grad.data) / world)
            off += p.numel()
        gns.append(nn.utils.clip_grad_norm_(agent.parameters(), args.max_grad_norm).detach())
        opt.step()
    return stats_all, gns, total / world


def dp_world_math(local_num_envs, num_steps, num_minibatches, world_size):
    """cleanrl/ppo_atari_multigpu.py:166-172."""
    local_batch_size = int(local_num_envs * num_steps)
    local_minibatch_size = int(local_batch_size // num_minibatches)
    num_envs = local_num_envs * world_size
    batch_size = int(num_envs * num_steps)
    minibatch_size = int(batch_size // num_minibatches)
    return dict(local_batch_size=local_batch_size, local_minibatch_size=local_minibatch_size,
                num_envs=num_envs, batch_size=batch_size, minibatch_size=minibatch_size)
