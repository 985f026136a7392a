"""Host env feed: batched object-vector assembly and reward normalisation in native host code.

The reference builds every observation in Python, per env and per step -- ``OCAtari.ns_state`` chains the ``[x, y]`` of
all object slots into a list (``submodules/OC_Atari/ocatari/core.py:444-448``), a ``deque(maxlen=4)`` stacks the frames
(``core.py:152-153,274-283``), ``np.array(deque)`` converts it (``core.py:213-214``), ``SubprocVecEnv`` stacks the N
arrays and ``VecNormalize(norm_obs=False, norm_reward=True)`` rescales the rewards (``cleanrl/ppo_atari_oc.py:411-414``)
-- and then creates three pageable tensors per step (``:512-514``).  ``ObjectFeed`` takes the object slots of all N
envs of one step (what an env worker reads out of the emulator's RAM: position and visibility per slot) and writes the
``(N, frames, 2 * slots)`` fp32 block directly into a caller-owned buffer -- the rollout collector's pinned staging
memory -- through ``libocppo.so`` (``ocppo_feed_*``, ``include/ocppo.h``).  ALE and OCAtari are not installed on this
box; the emulator stays with the env workers either way (host CPU).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi


class ObjectFeed:
    def __init__(self, num_envs: int, n_slots: int, frames: int = 4, gamma: float = 0.99, epsilon: float = 1e-8,
                 clip_reward: float = 10.0):
        self.N, self.S, self.F = int(num_envs), int(n_slots), int(frames)
        self._f = C.c_void_p()
        _abi.check(_abi.lib().ocppo_feed_create(self.N, self.F, self.S, float(gamma), float(epsilon), float(clip_reward),
                                                C.byref(self._f)))

    @property
    def obs_shape(self):
        return (self.N, self.F, 2 * self.S)

    @staticmethod
    def _p(a):
        return C.c_void_p(a.ctypes.data)

    def push(self, xy, visible, refill=None, out=None):
        """xy ``(N, slots, 2)`` integer positions, visible ``(N, slots)``, refill ``(N,)`` (True where the env was reset:
        the whole frame stack becomes this state) -> ``out`` fp32 ``(N, frames, 2*slots)`` (allocated if not given; pass
        the numpy view of a pinned tensor to have the block written in place)."""
        xy = np.ascontiguousarray(xy, dtype=np.int32)
        visible = np.ascontiguousarray(visible, dtype=np.uint8)
        if xy.shape != (self.N, self.S, 2) or visible.shape != (self.N, self.S):
            raise _abi.OcppoError(f"expected xy {(self.N, self.S, 2)} and visible {(self.N, self.S)}, got {xy.shape} / {visible.shape}")
        rf = None
        if refill is not None:
            rf = np.ascontiguousarray(refill, dtype=np.uint8)
            if rf.shape != (self.N,):
                raise _abi.OcppoError("refill must have one flag per env")
        if out is None:
            out = np.empty(self.obs_shape, dtype=np.float32)
        if out.dtype != np.float32 or out.size != self.N * self.F * 2 * self.S or not out.flags.c_contiguous:
            raise _abi.OcppoError("out must be a C-contiguous float32 array of (N, frames, 2*slots) elements")
        _abi.check(_abi.lib().ocppo_feed_push(self._f, self._p(xy), self._p(visible), None if rf is None else self._p(rf),
                                              self._p(out)))
        return out

    def normalize_reward(self, reward, done, out_reward=None, out_done=None):
        """``VecNormalize.step_wait`` reward path: -> (fp32 normalised rewards, fp32 done flags)."""
        reward = np.ascontiguousarray(reward, dtype=np.float64)
        done = np.ascontiguousarray(done, dtype=np.uint8)
        if reward.shape != (self.N,) or done.shape != (self.N,):
            raise _abi.OcppoError("reward / done must have one entry per env")
        if out_reward is None:
            out_reward = np.empty(self.N, dtype=np.float32)
        if out_done is None:
            out_done = np.empty(self.N, dtype=np.float32)
        for o in (out_reward, out_done):
            if o.dtype != np.float32 or o.size != self.N or not o.flags.c_contiguous:
                raise _abi.OcppoError("outputs must be C-contiguous float32 arrays of N elements")
        _abi.check(_abi.lib().ocppo_feed_normalize_reward(self._f, self._p(reward), self._p(done), self._p(out_reward),
                                                          self._p(out_done)))
        return out_reward, out_done

    def state(self):
        """-> (ret_rms mean, var, count, per-env discounted returns)."""
        mvc = np.empty(3, dtype=np.float64)
        ret = np.empty(self.N, dtype=np.float64)
        _abi.check(_abi.lib().ocppo_feed_state(self._f, self._p(mvc), self._p(ret)))
        return float(mvc[0]), float(mvc[1]), float(mvc[2]), ret

    def close(self):
        if self._f:
            _abi.lib().ocppo_feed_destroy(self._f)
            self._f = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
