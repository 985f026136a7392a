"""B200-native PPO_OBJ learner: drop-in for the object-centric path of
BluemlJ/oc_cleanrl's ``cleanrl/ppo_atari_oc.py`` (``--obs_mode obj
--architecture PPO_OBJ``).  All compute is in ``csrc/libocppo.so`` (hand-written
sm_100a CUDA behind the C ABI of ``include/ocppo.h``); there is no fallback."""

__version__ = "0.1.0"
