"""The oracle against the reference's golden vectors / known answers (CPU only)."""
import ctypes

import numpy as np
import pytest
import torch

from conftest import REFERENCE_PRESENT, load_golden
from oracle import ref_ppo, ref_sample
from oc_cleanrl_b200.synthetic import make_gae_inputs, make_rollout

STAT_KEYS = ("pg_loss", "v_loss", "entropy", "loss", "old_approx_kl", "approx_kl", "clipfrac")


def _c_gae(lib, r, v, d, nv, nd, gamma, lam):
    T, N = r.shape
    arrs = [np.ascontiguousarray(x, dtype=np.float32) for x in (r, v, d, nv, nd)]
    adv = np.zeros((T, N), np.float32)
    ret = np.zeros((T, N), np.float32)
    P = lambda x: x.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    rc = lib.ocppo_oracle_gae_f32(*[P(x) for x in arrs], ctypes.c_int64(T), ctypes.c_int64(N),
                                  ctypes.c_double(gamma), ctypes.c_double(lam), P(adv), P(ret))
    assert rc == 0
    return adv, ret


# shapes of the reference's only numeric test (tests/test_jax_compute_gae.py:12-18) plus edge cases
@pytest.mark.parametrize("T,N", [(123, 7), (128, 512), (1, 1), (2, 33), (5, 4)])
def test_gae_three_statements_agree_bitwise(oracle_lib, T, N):
    d = make_gae_inputs(T, N, seed=42)
    adv, ret = ref_ppo.ref_gae(d["rewards"], d["values"], d["dones"], d["next_value"], d["next_done"], 0.99, 0.95)
    a2, r2 = ref_ppo.np_gae_fp32(*[d[k].numpy() for k in ("rewards", "values", "dones", "next_value", "next_done")],
                                 0.99, 0.95)
    a3, r3 = _c_gae(oracle_lib, *[d[k].numpy() for k in ("rewards", "values", "dones", "next_value", "next_done")],
                    0.99, 0.95)
    assert np.array_equal(adv.numpy(), a2) and np.array_equal(ret.numpy(), r2)
    assert np.array_equal(adv.numpy(), a3) and np.array_equal(ret.numpy(), r3)


def test_gae_done_masks_bootstrap():
    T, N = 4, 2
    r = torch.ones(T, N)
    v = torch.zeros(T, N)
    d = torch.zeros(T, N)
    d[2, 0] = 1.0                                # episode boundary before step 2 of env 0
    adv, _ = ref_ppo.ref_gae(r, v, d, torch.zeros(N), torch.zeros(N), 0.5, 1.0)
    assert adv[1, 0].item() == 1.0               # no bootstrap across the done
    assert adv[1, 1].item() == pytest.approx(1.0 + 0.5 * (1.0 + 0.5 * 1.0))


def test_numpy_shuffle_known_answer():
    g = load_golden("numpy_shuffle_seed42")
    rs = np.random.RandomState(42)
    b = np.arange(524288)
    rs.shuffle(b)
    assert list(b[:5]) == [293397, 493367, 445980, 409194, 406340]
    assert np.array_equal(b[:16], g["first16"]) and np.array_equal(b[-16:], g["last16"])
    chk = np.int64((b.astype(np.int64) * np.arange(524288, dtype=np.int64)).sum() % (2**61 - 1))
    assert chk == g["checksum"]
    assert np.array_equal(np.sort(b), np.arange(524288))


def test_philox_known_answer_vectors():
    # Random123 kat_vectors, philox4x32-10
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
        ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
        ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
         [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
    ]
    for ctr, key, want in kat:
        got = ref_sample.philox4x32_10(np.array([ctr], np.uint32), np.array([key], np.uint32))[0]
        assert [int(x) for x in got] == want


def test_multinomial_single_draw_is_exponential_race():
    """torch.multinomial(p, 1, True) == argmax(p / Exp(1)) drawn from the same generator
    (categorical.py:144-149 -> ATen multinomial fast path); SURVEY.md A.3."""
    p = torch.softmax(torch.randn(1000, 6, generator=torch.Generator().manual_seed(0)), -1)
    g1 = torch.Generator().manual_seed(5)
    g2 = torch.Generator().manual_seed(5)
    s = torch.multinomial(p, 1, True, generator=g1).squeeze(-1)
    q = torch.empty_like(p).exponential_(1, generator=g2)
    assert torch.equal(s, torch.argmax(p / q, dim=-1))


def test_sampler_layout_and_statistics():
    numel = 4096 * 6
    grid, inc = ref_sample.torch_cuda_grid(numel)
    assert (grid, inc) == (96, 4)
    assert ref_sample.torch_cuda_grid(148 * 8 * 256 * 4 + 1) == (1184, 8)
    z = np.log(np.array([[0.5, 0.25, 0.125, 0.125]] * 20000, np.float32))
    out = ref_sample.ref_categorical_sample(z, seed=42, offset=0)
    freq = np.bincount(out["action"], minlength=4) / 20000.0
    assert np.abs(freq - np.array([0.5, 0.25, 0.125, 0.125])).max() < 0.015
    assert np.allclose(out["logp"], np.log([0.5, 0.25, 0.125, 0.125])[out["action"]], atol=1e-6)
    # offset moves the stream; same (seed, offset) reproduces it
    again = ref_sample.ref_categorical_sample(z, seed=42, offset=0)
    moved = ref_sample.ref_categorical_sample(z, seed=42, offset=4)
    assert np.array_equal(out["action"], again["action"]) and not np.array_equal(out["action"], moved["action"])


@pytest.mark.parametrize("name", ["small_pong", "small_wide", "default_pong"])
def test_oracle_reproduces_reference_golden(name):
    """OraclePPObj + restated loop vs fixtures produced with the reference's own PPObj."""
    g = load_golden(name)
    torch.set_num_threads(1)
    F, D, A, T, N = (int(g[k]) for k in ("F", "D", "A", "T", "N"))
    enc, dec = tuple(int(x) for x in g["enc"]), tuple(int(x) for x in g["dec"])
    args = ref_ppo.PPOArgs(num_envs=N, num_steps=T, update_epochs=int(g["epochs"]),
                           num_minibatches=int(g["minibatches"]), encoder_dims=enc, decoder_dims=dec)
    agent = ref_ppo.make_agent(F, D, A, enc, dec, seed=42)
    init = ref_ppo.flat_params(agent)
    assert float(init.double().sum()) == pytest.approx(float(g["init_sum"]), rel=1e-9)
    data = make_rollout(T, N, F, D, seed=1234)
    actions = torch.as_tensor(g["actions"])
    with torch.no_grad():
        _, logp, ent, value = agent.get_action_and_value(data["obs"].reshape(T * N, F, D), actions.reshape(-1))
    assert np.allclose(logp.numpy(), g["logp"], rtol=1e-5, atol=1e-6)
    assert np.allclose(ent.numpy(), g["entropy"], rtol=1e-5, atol=1e-6)
    assert np.allclose(value.numpy().reshape(-1), g["value"], rtol=1e-5, atol=1e-5)
    values_old, logprobs_old = torch.as_tensor(g["values_old"]), torch.as_tensor(g["logprobs_old"])
    adv, ret = ref_ppo.ref_gae(data["rewards"], values_old, data["dones"], torch.as_tensor(g["next_value"]),
                               data["next_done"], args.gamma, args.gae_lambda)
    assert np.array_equal(adv.numpy(), g["adv"]) and np.array_equal(ret.numpy(), g["ret"])
    opt = ref_ppo.make_optimizer(agent, args)
    trace = []
    ref_ppo.ref_update(agent, opt, data["obs"], actions, logprobs_old, adv, ret, values_old, args,
                       np_rng=np.random.RandomState(7), trace=trace)
    got = np.array([[float(t["stats"][k]) for k in STAT_KEYS] for t in trace])
    assert np.allclose(got, g["trace_stats"], rtol=2e-4, atol=1e-6)
    assert np.allclose([float(t["gn"]) for t in trace], g["trace_gn"], rtol=2e-4)
    assert np.array_equal(trace[0]["mb_inds"], g["mb0"])
    final = ref_ppo.flat_params(agent).numpy()
    if "final_flat" in g:
        assert np.abs(final - g["final_flat"]).max() <= 1e-4 * np.abs(g["final_flat"]).max()
    else:
        assert np.abs(final[::997] - g["final_sample"]).max() <= 1e-4 * np.abs(g["final_sample"]).max()


@pytest.mark.skipif(not REFERENCE_PRESENT, reason="/root/reference only exists in the build container")
def test_oracle_model_equals_reference_model():
    for (F, D, A, enc, dec) in [(4, 6, 6, (256, 512, 1024, 512), (512,)), (4, 88, 6, (32, 16), (8,)),
                                (2, 5, 3, (7,), (9, 4))]:
        ref = ref_ppo.make_agent(F, D, A, enc, dec, seed=42, use_reference=True)
        ours = ref_ppo.make_agent(F, D, A, enc, dec, seed=42)
        assert list(ref.state_dict().keys()) == list(ours.state_dict().keys())
        for (k, a), (_, b) in zip(ref.state_dict().items(), ours.state_dict().items()):
            assert torch.equal(a, b), k
        x = torch.randn(9, F, D)
        act = torch.randint(0, A, (9,))
        with torch.no_grad():
            for u, v in zip(ref.get_action_and_value(x, act), ours.get_action_and_value(x, act)):
                assert torch.equal(u, v)
            assert torch.equal(ref.get_value(x), ours.get_value(x))


def test_relu_decisions_context_reproduces_and_overrides():
    """oracle/ref_ppo.py::relu_decisions (the hook the GPU parity test uses to hand the oracle the device path's ReLU
    decisions): with the network's OWN decisions the loss and every gradient are bit-identical to the plain
    evaluation; flipping one decision of a unit whose pre-activation is far from zero changes that unit's row."""
    torch.manual_seed(0)
    F, D, A = 4, 6, 6
    agent = ref_ppo.make_agent(F, D, A, (32, 64), (32,), seed=3)
    rs = np.random.RandomState(0)
    n = 64
    batch = (torch.as_tensor(rs.randint(0, 200, (n, F, D)), dtype=torch.float32), torch.as_tensor(rs.randint(0, A, n)),
             torch.as_tensor(-1.79 + 0.05 * rs.randn(n), dtype=torch.float32),
             torch.as_tensor(rs.randn(n), dtype=torch.float32), torch.as_tensor(rs.randn(n), dtype=torch.float32),
             torch.as_tensor(rs.randn(n), dtype=torch.float32))
    idx = np.arange(n)
    loss, _ = ref_ppo.ref_minibatch_loss(agent, *batch, idx, ref_ppo.PPOArgs())
    loss.backward()
    plain = ref_ppo.flat_grads(agent).clone()
    with ref_ppo.relu_decisions(agent) as rd:
        ref_ppo.ref_minibatch_loss(agent, *batch, idx, ref_ppo.PPOArgs())
    own = [r["decision"].clone() for r in rd.records]
    assert len(own) == 3 and own[0].shape == (n, F, 32) and own[-1].shape == (n, 32)
    for p in agent.parameters():
        p.grad = None
    with ref_ppo.relu_decisions(agent, own):
        loss2, _ = ref_ppo.ref_minibatch_loss(agent, *batch, idx, ref_ppo.PPOArgs())
    loss2.backward()
    assert torch.equal(loss2, loss) and torch.equal(ref_ppo.flat_grads(agent), plain)
    # switch off every decision of unit 5 of the last trunk layer: only then does its weight-gradient row vanish
    flipped = [m.clone() for m in own]
    flipped[-1][:, 5] = False
    for p in agent.parameters():
        p.grad = None
    with ref_ppo.relu_decisions(agent, flipped):
        loss3, _ = ref_ppo.ref_minibatch_loss(agent, *batch, idx, ref_ppo.PPOArgs())
    loss3.backward()
    last_linear = [m for m in agent.network if isinstance(m, torch.nn.Linear)][-1]
    assert torch.count_nonzero(last_linear.weight.grad[5]) == 0
    if own[-1][:, 5].any():
        assert not torch.equal(ref_ppo.flat_grads(agent), plain)
