"""CPU restatement of the reference's host-side observation assembly and reward normalisation.

TEST INFRASTRUCTURE -- see ``oracle/__init__.py`` for who may import this.

* ``RefObjectEnv`` follows OCAtari's object-mode observation path statement by statement with the same containers
  (python lists, ``itertools.chain``, ``collections.deque``): ``submodules/OC_Atari/ocatari/core.py:444-448``
  (``ns_state``), ``ocatari/ram/game_objects.py:184-188`` (``_nsrepr``), ``core.py:152-153`` (the deque),
  ``core.py:274-283`` (``_fill_buffer``), ``core.py:241-244`` (``_reset_buffer``), ``core.py:213-214,270-271``
  (``np.array(deque)``).  The emulator is replaced by a stub that is handed the object slots.
* ``RefVecNormalizeReward`` restates ``VecNormalize(venv, norm_obs=False, norm_reward=True)``
  (``cleanrl/ppo_atari_oc.py:411-414``) of the pinned third-party dependency **stable-baselines3==2.0.0**
  (``requirements/requirements.txt``; not vendored under ``/root/reference`` and not installed here):
  ``common/running_mean_std.py`` (``RunningMeanStd.update`` / ``update_from_moments``, Chan et al.'s parallel variance)
  and ``common/vec_env/vec_normalize.py`` (``step_wait``, ``_update_reward``, ``normalize_reward``), with numpy
  executing the arithmetic as there.  **Parity unpinned**: the reference ships no vectors for this component and
  neither OCAtari's emulator side nor SB3 can be imported here.
"""
from __future__ import annotations

from collections import deque
from itertools import chain

import numpy as np


class _Slot:
    """ocatari/ram/game_objects.py: the three attributes ``_nsrepr`` reads."""

    def __init__(self):
        self.x, self.y, self._visible = 0, 0, False

    @property
    def _nsrepr(self):                               # game_objects.py:184-188
        if not self._visible:
            return [0, 0]
        return [self.x, self.y]


class RefObjectEnv:
    """One env's observation path in ``obs_mode="obj"``."""

    def __init__(self, n_slots, buffer_window_size=4):
        self.objects = [_Slot() for _ in range(n_slots)]                       # core.py:122-124
        self.buffer_window_size = buffer_window_size
        self._state_buffer_ns = deque([], maxlen=buffer_window_size)           # core.py:152-153

    @property
    def ns_state(self):                              # core.py:444-448
        return list(chain.from_iterable([o._nsrepr for o in self.objects]))

    def _detect(self, xy, visible):                  # stands for detect_objects(): RAM -> slots
        for o, (x, y), v in zip(self.objects, xy, visible):
            o.x, o.y, o._visible = int(x), int(y), bool(v)

    def _fill_buffer(self):                          # core.py:274-283
        self._state_buffer_ns.append(self.ns_state)

    def reset(self, xy, visible):                    # core.py:260-271
        self._detect(xy, visible)
        for _ in range(self.buffer_window_size):     # core.py:241-244
            self._fill_buffer()
        return np.array(self._state_buffer_ns)

    def step(self, xy, visible):                     # core.py:205-214
        self._detect(xy, visible)
        self._fill_buffer()
        return np.array(self._state_buffer_ns)


class RefRunningMeanStd:
    """stable_baselines3/common/running_mean_std.py (2.0.0)."""

    def __init__(self, epsilon=1e-4, shape=()):
        self.mean = np.zeros(shape, np.float64)
        self.var = np.ones(shape, np.float64)
        self.count = epsilon

    def update(self, arr):
        batch_mean = np.mean(arr, axis=0)
        batch_var = np.var(arr, axis=0)
        batch_count = arr.shape[0]
        self.update_from_moments(batch_mean, batch_var, batch_count)

    def update_from_moments(self, batch_mean, batch_var, batch_count):
        delta = batch_mean - self.mean
        tot_count = self.count + batch_count
        new_mean = self.mean + delta * batch_count / tot_count
        m_a = self.var * self.count
        m_b = batch_var * batch_count
        m_2 = m_a + m_b + np.square(delta) * self.count * batch_count / (self.count + batch_count)
        new_var = m_2 / (self.count + batch_count)
        new_count = batch_count + self.count
        self.mean = new_mean
        self.var = new_var
        self.count = new_count


class RefVecNormalizeReward:
    """stable_baselines3/common/vec_env/vec_normalize.py (2.0.0), reward path only (``norm_obs=False``)."""

    def __init__(self, num_envs, gamma=0.99, epsilon=1e-8, clip_reward=10.0):
        self.ret_rms = RefRunningMeanStd(shape=())
        self.returns = np.zeros(num_envs)
        self.gamma, self.epsilon, self.clip_reward = gamma, epsilon, clip_reward

    def step_wait(self, rewards, dones):
        self.returns = self.returns * self.gamma + rewards            # _update_reward
        self.ret_rms.update(self.returns)
        rewards = np.clip(rewards / np.sqrt(self.ret_rms.var + self.epsilon), -self.clip_reward, self.clip_reward)
        self.returns[dones] = 0
        return rewards
