"""Mirror of the reference's ``cleanrl/architectures/common.py:7-16``."""
import numpy as np
import torch
import torch.nn as nn


def layer_init(layer, std=np.sqrt(2), bias_const=0.0):
    """Orthogonal weight / constant bias (common.py:7-10)."""
    torch.nn.init.orthogonal_(layer.weight, std)
    torch.nn.init.constant_(layer.bias, bias_const)
    return layer


class Predictor(nn.Module):
    """``predict`` = greedy action (common.py:13-16); needs .network/.actor/.device."""

    def predict(self, x, states=None, **_):
        with torch.no_grad():
            logits = self.actor(self.network(torch.Tensor(x).to(self.device)))
            return np.argmax(logits.cpu().numpy(), axis=1), states
