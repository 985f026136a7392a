"""Host env feed against the numpy / pure-python restatement of the reference's per-env observation path
(oracle/ref_feed.py: OCAtari core.py:444-448 + deque semantics) and of SB3 2.0.0's VecNormalize reward path.
A stub object source stands for ALE's RAM (unavailable on this box).  Bit-exact: the block is integers in fp32, the
running statistics follow numpy's own reduction order."""
import numpy as np
import pytest

from oracle import ref_feed


def _stream(rng, N, S, steps, p_done=0.05, p_invisible=0.2):
    for _ in range(steps):
        xy = np.stack([rng.randint(0, 160, (N, S)), rng.randint(0, 210, (N, S))], axis=-1)
        yield xy, rng.rand(N, S) >= p_invisible, rng.rand(N) < p_done


@pytest.mark.parametrize("N,S,F", [(1, 3, 4), (7, 3, 4), (16, 44, 4), (5, 110, 4), (3, 3, 1), (4, 2, 6)])
def test_frame_stacked_object_vectors_equal_the_reference_path(built_lib, N, S, F):
    from oc_cleanrl_b200.feed import ObjectFeed
    rng = np.random.RandomState(N * 100 + S)
    feed = ObjectFeed(N, S, F)
    envs = [ref_feed.RefObjectEnv(S, F) for _ in range(N)]
    out = np.full((N, F, 2 * S), -1.0, dtype=np.float32)
    first = True
    for xy, vis, done in _stream(rng, N, S, 40):
        # SubprocVecEnv auto-resets an env whose episode ended: its next observation is the reset observation
        refill = np.ones(N, bool) if first else done
        want = np.stack([e.reset(xy[n], vis[n]) if refill[n] else e.step(xy[n], vis[n]) for n, e in enumerate(envs)])
        got = feed.push(xy, vis, refill, out=out)
        assert got is out
        # torch.tensor(next_obs_np, dtype=torch.float32) of ppo_atari_oc.py:513
        assert np.array_equal(got, want.astype(np.float32))
        assert want.shape == (N, F, 2 * S)
        first = False


def test_first_push_fills_the_stack_without_a_refill_flag(built_lib):
    from oc_cleanrl_b200.feed import ObjectFeed
    feed = ObjectFeed(2, 3)
    xy = np.arange(12).reshape(2, 3, 2)
    vis = np.array([[1, 0, 1], [1, 1, 1]], bool)
    got = feed.push(xy, vis)
    e = ref_feed.RefObjectEnv(3)
    assert np.array_equal(got[0], e.reset(xy[0], vis[0]).astype(np.float32))
    assert (got[0, :, 2:4] == 0).all()          # the invisible slot is [0, 0] in every frame (game_objects.py:186-187)
    got2 = feed.push(xy + 1, vis)
    assert np.array_equal(got2[0], e.step(xy[0] + 1, vis[0]).astype(np.float32))


@pytest.mark.parametrize("N", [1, 5, 8, 100, 129, 4096])
def test_reward_normaliser_equals_vecnormalize_restatement(built_lib, N):
    from oc_cleanrl_b200.feed import ObjectFeed
    rng = np.random.RandomState(N)
    feed = ObjectFeed(N, 3)
    ref = ref_feed.RefVecNormalizeReward(N)
    for step in range(60):
        # Atari rewards: sparse, integer-valued floats; sometimes large
        reward = (rng.rand(N) < 0.1) * rng.choice([1.0, -1.0, 10.0, 200.0], N)
        done = rng.rand(N) < 0.02
        want = ref.step_wait(reward.copy(), done)
        got, got_done = feed.normalize_reward(reward, done)
        # ppo_atari_oc.py:512,514: float32 tensors of the normalised reward and of the done flags
        assert np.array_equal(got, want.astype(np.float32)), step
        assert np.array_equal(got_done, done.astype(np.float32))
        mean, var, count, returns = feed.state()
        assert mean == float(ref.ret_rms.mean) and var == float(ref.ret_rms.var) and count == float(ref.ret_rms.count)
        assert np.array_equal(returns, ref.returns)
    assert np.abs(got).max() <= 10.0


def test_shape_errors_are_raised(built_lib):
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200.feed import ObjectFeed
    feed = ObjectFeed(4, 3)
    with pytest.raises(_abi.OcppoError):
        feed.push(np.zeros((4, 2, 2)), np.zeros((4, 3)))
    with pytest.raises(_abi.OcppoError):
        feed.push(np.zeros((4, 3, 2)), np.zeros((4, 3)), out=np.zeros((4, 4, 6), np.float64))
    with pytest.raises(_abi.OcppoError):
        feed.normalize_reward(np.zeros(3), np.zeros(4))
    with pytest.raises(_abi.OcppoError):
        ObjectFeed(0, 3)
