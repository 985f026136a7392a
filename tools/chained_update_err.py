"""How far do 16 chained minibatch updates (4 epochs x 4 minibatches, Adam) drift?  8 envs x 128 steps (BASELINE
configs[0] size), default widths: post-update weights of this library (per GEMM engine) against the oracle loop in torch
fp32 and fp64 on the CPU, next to the oracle's own fp32-vs-fp64 distance -- the amplification any fp32 evaluation of
this loop shows (Adam divides every gradient by its running magnitude, so parameters whose gradient is near zero move by
+-lr whatever the size of the difference).   python tools/chained_update_err.py [engines...]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner, compute_gae
from oc_cleanrl_b200.synthetic import make_rollout, stub_envs
from oracle import ref_ppo

dev = torch.device("cuda:0")
lib = _abi.lib()
F, D, A, T, N = 4, 6, 6, 128, 8
enc, dec = (256, 512, 1024, 512), (512,)
data = make_rollout(T, N, F, D, seed=1234)
rel = lambda a, b: float(np.abs(a - b).max() / np.abs(b).max())

def rollout():
    assert lib.ocppo_set_gemm_mode(0) == 0
    torch.manual_seed(42)
    agent = PPObj(stub_envs(F, D, A), dev, enc, dec).to(dev)
    obs = data["obs"].to(dev)
    torch.cuda.manual_seed(0)
    with torch.no_grad():
        action, logp, _, value = agent.get_action_and_value(obs.reshape(T * N, F, D))
        next_value = agent.get_value(data["next_obs"].to(dev)).reshape(-1)
    values = value.reshape(T, N)
    adv, ret = compute_gae(data["rewards"].to(dev), values, data["dones"].to(dev), next_value, data["next_done"].to(dev), 0.99, 0.95)
    return tuple(t.cpu() for t in (action, logp, adv, ret, values))

def ours(engine, roll):
    assert lib.ocppo_set_gemm_mode(engine) == 0
    action, logp, adv, ret, values = (t.to(dev) for t in roll)
    torch.manual_seed(42)
    agent = PPObj(stub_envs(F, D, A), dev, enc, dec).to(dev)
    learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T), np_rng=np.random.RandomState(11))
    learner.update(data["obs"].to(dev), action.reshape(T, N), logp.reshape(T, N), adv, ret, values)
    return torch.cat([p.detach().reshape(-1) for p in agent.parameters()]).double().cpu().numpy()

def oracle(dtype, roll):
    action, logp, adv, ret, values = roll
    o = ref_ppo.make_agent(F, D, A, enc, dec, seed=42, dtype=dtype)
    oargs = ref_ppo.PPOArgs(num_envs=N, num_steps=T)
    cast = lambda t: t.to(dtype) if t.is_floating_point() else t
    ref_ppo.ref_update(o, ref_ppo.make_optimizer(o, oargs), cast(data["obs"]), action.reshape(T, N), cast(logp).reshape(T, N), cast(adv), cast(ret), cast(values), oargs, np_rng=np.random.RandomState(11))
    return ref_ppo.flat_params(o).double().numpy()

roll = rollout()
o32, o64 = oracle(torch.float32, roll), oracle(torch.float64, roll)
print(f"oracle fp32 vs oracle fp64 after 16 chained steps: {rel(o32, o64):.2e}")
for e in [int(x) for x in sys.argv[1:]] or [0, 6, 8]:
    w = ours(e, roll)
    print(f"engine {e}: vs oracle fp32 {rel(w, o32):.2e}  vs oracle fp64 {rel(w, o64):.2e}", flush=True)
