"""Writes tests/golden/*.npz from the REFERENCE's own PPObj class.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
The learner loop of the reference is un-importable ``__main__`` code, so the loop
around the reference model is the operator-for-operator restatement in
``oracle/ref_ppo.py`` (GAE :533-547, loss :566-602, clip/Adam :605-610); the
model class itself (architectures/ppo.py:60-95) is imported unmodified.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_ppo  # noqa: E402
from oc_cleanrl_b200.synthetic import make_rollout  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def case(name, F, D, A, enc, dec, T, N, epochs, minibatches, store_full):
    torch.set_num_threads(1)            # fixed reduction order for reproducible fixtures
    agent = ref_ppo.make_agent(F, D, A, enc, dec, seed=42, use_reference=True)
    args = ref_ppo.PPOArgs(num_envs=N, num_steps=T, update_epochs=epochs, num_minibatches=minibatches,
                           encoder_dims=enc, decoder_dims=dec)
    data = make_rollout(T, N, F, D, seed=1234)
    rs = np.random.RandomState(99)
    actions = torch.as_tensor(rs.randint(0, A, size=(T, N)), dtype=torch.long)
    flat_obs = data["obs"].reshape(T * N, F, D)
    with torch.no_grad():
        _, logp, ent, value = agent.get_action_and_value(flat_obs, actions.reshape(-1))
        next_value = agent.get_value(data["next_obs"]).reshape(-1)
        hidden = agent.network(flat_obs)
        logits = agent.actor(hidden)
    values = value.reshape(T, N)
    logprobs = logp.reshape(T, N)
    # perturb old logprobs/values a little so ratios are not exactly 1 and clipping is exercised
    logprobs = logprobs + 0.15 * torch.as_tensor(rs.randn(T, N), dtype=torch.float32)
    values_old = values + 0.2 * torch.as_tensor(rs.randn(T, N), dtype=torch.float32)
    adv, ret = ref_ppo.ref_gae(data["rewards"], values_old, data["dones"], next_value, data["next_done"],
                               args.gamma, args.gae_lambda)
    init_flat = ref_ppo.flat_params(agent).clone()
    opt = ref_ppo.make_optimizer(agent, args)
    trace = []
    # first minibatch gradient (before clip), separately, on a twin agent
    twin = ref_ppo.make_agent(F, D, A, enc, dec, seed=42, use_reference=True)
    perm_rng = np.random.RandomState(7)
    b_inds = np.arange(T * N)
    perm_rng.shuffle(b_inds)
    mb0 = b_inds[: (T * N) // minibatches].copy()
    batch = (flat_obs, actions.reshape(-1), logprobs.reshape(-1), adv.reshape(-1), ret.reshape(-1),
             values_old.reshape(-1))
    loss0, stats0 = ref_ppo.ref_minibatch_loss(twin, *batch, mb0, args)
    loss0.backward()
    grad0 = ref_ppo.flat_grads(twin)
    stats, clipfracs, gn = ref_ppo.ref_update(agent, opt, data["obs"], actions, logprobs, adv, ret, values_old, args,
                                              np_rng=np.random.RandomState(7), trace=trace)
    final_flat = ref_ppo.flat_params(agent)
    out = dict(
        F=F, D=D, A=A, enc=np.array(enc), dec=np.array(dec), T=T, N=N, epochs=epochs, minibatches=minibatches,
        actions=actions.numpy(), logp=logp.numpy(), entropy=ent.numpy(), value=value.numpy().reshape(-1),
        next_value=next_value.numpy(), logits=logits.numpy(),
        logprobs_old=logprobs.numpy(), values_old=values_old.numpy(), adv=adv.numpy(), ret=ret.numpy(),
        mb0=mb0, loss0=float(loss0), grad0_norm=float(grad0.norm()),
        stats0=np.array([float(stats0[k]) for k in ("pg_loss", "v_loss", "entropy", "loss", "old_approx_kl",
                                                     "approx_kl", "clipfrac")]),
        trace_stats=np.array([[float(t["stats"][k]) for k in ("pg_loss", "v_loss", "entropy", "loss",
                                                               "old_approx_kl", "approx_kl", "clipfrac")]
                              for t in trace]),
        trace_gn=np.array([float(t["gn"]) for t in trace]),
        init_sum=float(init_flat.double().sum()), init_abs=float(init_flat.double().abs().sum()),
        final_sum=float(final_flat.double().sum()),
    )
    per_tensor = []
    off = 0
    for p in twin.parameters():
        g = grad0[off:off + p.numel()]
        per_tensor.append([float(g.norm()), float(g.double().sum())])
        off += p.numel()
    out["grad0_per_tensor"] = np.array(per_tensor)
    if store_full:
        out.update(grad0=grad0.numpy(), init_flat=init_flat.numpy(), final_flat=final_flat.numpy())
    else:
        sl = slice(0, None, 997)            # strided sample of the 2.2 M-element vectors
        out.update(grad0_sample=grad0.numpy()[sl], init_sample=init_flat.numpy()[sl],
                   final_sample=final_flat.numpy()[sl])
    path = os.path.join(OUT, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(name, os.path.getsize(path) // 1024, "KiB; loss0", float(loss0), "gn", out["trace_gn"][:3])


if __name__ == "__main__":
    if ref_ppo.import_reference_ppobj() is None:
        sys.exit("needs /root/reference")
    case("small_pong", 4, 6, 6, (32, 64), (32,), T=8, N=8, epochs=2, minibatches=2, store_full=True)
    case("small_wide", 4, 22, 4, (48,), (40, 24), T=6, N=10, epochs=1, minibatches=3, store_full=True)
    case("default_pong", 4, 6, 6, (256, 512, 1024, 512), (512,), T=8, N=16, epochs=1, minibatches=2,
         store_full=False)
    # known answers that need no model
    np.random.seed(42)
    b = np.arange(524288)
    np.random.shuffle(b)
    np.savez_compressed(os.path.join(OUT, "numpy_shuffle_seed42.npz"), first16=b[:16], last16=b[-16:],
                        checksum=np.int64((b.astype(np.int64) * np.arange(524288, dtype=np.int64)).sum() % (2**61 - 1)))
    print("shuffle first5", b[:5])
