"""``.cleanrl_model`` checkpoints in the reference's format.

``cleanrl/ppo_atari_oc.py:484-491`` writes ``{"model_weights": agent.state_dict(), "args": vars(args), "Timesteps":
iteration * batch_size}`` every ``checkpoint_interval`` iterations and ``:675-677`` the same without ``Timesteps`` at the
end; ``architectures/loading.py:111-122`` reads ``ckpt["model_weights"]`` back with ``load_state_dict``.  The drop-in
``PPObj`` keeps the reference's parameter names, so files written here load into the reference class and vice versa.
The reference saves no optimizer state; ``optimizer`` here is optional and ignored by the reference's loader.
"""
from __future__ import annotations

import torch


def save_checkpoint(path, agent, args=None, timesteps=None, learner=None):
    """Write a reference-format checkpoint.  Weights are detached CPU copies (the live parameters are views of the flat
    device arena).  With `learner`, Adam's moments and step counter ride along under ``"optimizer"``."""
    data = {"model_weights": {k: v.detach().cpu().clone() for k, v in agent.state_dict().items()},
            "args": dict(vars(args)) if args is not None and not isinstance(args, dict) else dict(args or {})}
    if timesteps is not None:
        data["Timesteps"] = int(timesteps)
    if learner is not None:
        data["optimizer"] = {"exp_avg": agent.exp_avg.detach().cpu().clone(),
                             "exp_avg_sq": agent.exp_avg_sq.detach().cpu().clone(),
                             "step": int(learner.adam_step), "lr": float(learner.lr)}
    torch.save(data, path)
    return data


def load_checkpoint(path, agent, learner=None, map_location="cpu"):
    """Load a checkpoint written by the reference script or by `save_checkpoint` into `agent` (in place: the parameters
    stay views of the device arena).  Returns the checkpoint dict."""
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    agent.load_state_dict(ckpt["model_weights"])
    opt = ckpt.get("optimizer")
    if learner is not None and opt is not None:
        agent.exp_avg.copy_(opt["exp_avg"])
        agent.exp_avg_sq.copy_(opt["exp_avg_sq"])
        learner.adam_step = int(opt["step"])
        learner.lr = float(opt["lr"])
    return ckpt
