"""One FULL-SIZE minibatch (131,072 samples of a 4096 x 128 rollout, default widths: BASELINE configs[1] / [2]) through
``ocppo_minibatch_fwd_bwd`` against the oracle's operator sequence (``oracle.ref_ppo.ref_minibatch_loss`` =
cleanrl/ppo_atari_oc.py:566-602 + ``loss.backward()`` :608) run by torch ON THE DEVICE in fp32 and fp64 -- the CPU would
take minutes per case.  Shared by tests/test_gpu_parity.py (asserts the bars) and tools/fullsize_sweep.py (writes the
per-tensor table under profiles/)."""
import ctypes as C

import numpy as np
import torch

from oracle import ref_ppo
from oc_cleanrl_b200.synthetic import GAME_SHAPES, make_rollout, stub_envs

F, T, N = 4, 128, 4096
ENC, DEC = (256, 512, 1024, 512), (512,)
MB = 131072


def perturb_like_training(params, seed):
    """Weights a few thousand updates into training are no longer orthogonal with zero biases: every tensor gets a
    seeded multiplicative / additive disturbance (the same one on both sides, applied to the CPU copy first)."""
    g = torch.Generator().manual_seed(1000 + seed)
    for p in params:
        with torch.no_grad():
            if p.dim() == 2:
                p.mul_(1.0 + 0.3 * torch.randn(p.shape, generator=g))
                p.add_(0.02 * p.abs().max() * torch.randn(p.shape, generator=g))
            else:
                p.add_(0.1 * torch.randn(p.shape, generator=g))


def permuted_twin(model, seed):
    """The SAME function with every hidden layer's units renumbered (rows of W_l and b_l permuted, columns of W_{l+1}
    permuted to match, frame blocks of the Flatten respected): torch evaluates it with the same kernels but sums every dot
    product in a different order -- an independent, equally valid fp32 evaluation of the reference's operator sequence.
    -> (twin model, function mapping the twin's per-tensor gradients back to the original numbering)."""
    import copy
    g = torch.Generator().manual_seed(7000 + seed)
    twin = copy.deepcopy(model)
    lins = [m for m in twin.network if isinstance(m, torch.nn.Linear)]
    n_enc = len(ENC)
    perms, col_perms = [], []
    prev = None
    with torch.no_grad():
        for l, lin in enumerate(lins):
            out_dim, in_dim = lin.weight.shape
            if prev is None:
                cp = torch.arange(in_dim)
            elif l == n_enc:                       # first decoder layer: input index = frame * width + unit
                width = prev.numel()
                cp = torch.cat([f * width + prev for f in range(in_dim // width)])
            else:
                cp = prev
            P = torch.randperm(out_dim, generator=g)
            lin.weight.copy_(lin.weight[P][:, cp])
            lin.bias.copy_(lin.bias[P])
            perms.append(P); col_perms.append(cp)
            prev = P
        for head in (twin.actor, twin.critic):
            head.weight.copy_(head.weight[:, prev])

    def unpermute(grads):
        out = []
        for l in range(len(lins)):
            gw = grads[2 * l].reshape(lins[l].weight.shape)
            inv_r, inv_c = np.argsort(perms[l].numpy()), np.argsort(col_perms[l].numpy())
            out.append(gw[inv_r][:, inv_c].reshape(-1))
            out.append(grads[2 * l + 1][inv_r])
        k = 2 * len(lins)
        inv_c = np.argsort(prev.numpy())
        out.append(grads[k].reshape(twin.actor.weight.shape)[:, inv_c].reshape(-1)); out.append(grads[k + 1])
        out.append(grads[k + 2].reshape(twin.critic.weight.shape)[:, inv_c].reshape(-1)); out.append(grads[k + 3])
        return out
    return twin, unpermute


def rel_max(got, want):
    d = np.abs(want).max()
    return float(np.abs(got - want).max() / (d if d > 0 else 1.0))


def rel_rows(got, want, shape):
    """Per-row figure for a weight tensor (out, in): max over rows of  max|delta row| / max|reference row| ."""
    if len(shape) != 2:
        return rel_max(got, want)
    g, w = got.reshape(shape), want.reshape(shape)
    den = np.abs(w).max(axis=1)
    num = np.abs(g - w).max(axis=1)
    ok = den > 0
    return float((num[ok] / den[ok]).max()) if ok.any() else 0.0


def fullsize_case(dev, game, seed, engine, trained_like=False, with_masks=False, with_twin=False):
    """-> dict(shapes, e32, e64, ref, rows32 [, e64m]): per parameter tensor, the max-norm relative error of our gradient
    against torch fp32 / fp64, torch fp32 against fp64, the per-row figure against torch fp32, and optionally against fp64
    evaluated under the device path's own ReLU decisions; ``twin32``: torch fp32 on the unit-permuted twin (same function,
    another summation order) against torch fp32 -- how far two valid fp32 evaluations of the reference are apart."""
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200.architectures.ppo import PPObj
    from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
    lib = _abi.lib()
    prev = lib.ocppo_get_gemm_mode()
    assert lib.ocppo_set_gemm_mode(engine) == 0
    try:
        D, A = GAME_SHAPES[game]
        n = T * N
        cpu_model = ref_ppo.make_agent(F, D, A, ENC, DEC, seed=42 + seed)
        if trained_like:
            perturb_like_training(list(cpu_model.parameters()), seed)
        torch.manual_seed(42 + seed)
        agent = PPObj(stub_envs(F, D, A), dev, ENC, DEC).to(dev)
        with torch.no_grad():
            for p, q in zip(agent.parameters(), cpu_model.parameters()):
                p.copy_(q.to(dev))
        data = make_rollout(T, N, F, D, seed=1234 + 17 * seed)
        rs = np.random.RandomState(1 + seed)
        gpu = (data["obs"].reshape(n, F, D).to(dev), torch.as_tensor(rs.randint(0, A, n)).to(dev),
               torch.as_tensor(-np.log(A) + 0.05 * rs.randn(n), dtype=torch.float32).to(dev),
               torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev), torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev),
               torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev))
        flat = (gpu[0].reshape(n, -1),) + gpu[1:]
        perm = torch.as_tensor(np.random.RandomState(42 + seed).permutation(n)[:MB], dtype=torch.int64).to(dev)
        learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
        agent._ensure_workspace(MB, training=True, exact=True)
        _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(perm), MB,
                                               C.byref(learner._hp), _abi.ptr(learner._stats[0]), agent._stream()))
        torch.cuda.synchronize()
        shapes = [tuple(p.shape) for p in agent.parameters()]
        ours = [agent._grad_view(i).reshape(-1).double().cpu().numpy() for i in range(len(shapes))]
        masks = None
        if with_masks:
            masks = []
            for l, width in enumerate(ENC + DEC):
                act = torch.empty(MB * (F if l < len(ENC) else 1), width, device=dev)
                _abi.check(lib.ocppo_read_activation(agent._handle.raw, l, MB, _abi.ptr(act), agent._stream()))
                torch.cuda.synchronize()
                masks.append(act > 0)
                del act
        del agent, learner
        torch.cuda.empty_cache()

        def torch_grads(dtype, m=None, source=None):
            o = ref_ppo.make_agent(F, D, A, ENC, DEC, seed=0, dtype=dtype).to(dev)
            with torch.no_grad():
                for p, q in zip(o.parameters(), (source or cpu_model).parameters()):
                    p.copy_(q.to(dev).to(dtype))
            args_t = tuple(t.to(dtype) if t.is_floating_point() else t for t in gpu)
            if m is None:
                l, _ = ref_ppo.ref_minibatch_loss(o, *args_t, perm, ref_ppo.PPOArgs())
            else:
                with ref_ppo.relu_decisions(o, m):
                    l, _ = ref_ppo.ref_minibatch_loss(o, *args_t, perm, ref_ppo.PPOArgs())
            l.backward()
            g = [q.grad.reshape(-1).double().cpu().numpy() for q in o.parameters()]
            del o, args_t, l
            torch.cuda.empty_cache()
            return g
        assert not torch.backends.cuda.matmul.allow_tf32
        g32, g64 = torch_grads(torch.float32), torch_grads(torch.float64)
        out = dict(shapes=shapes,
                   e32=[rel_max(a, b) for a, b in zip(ours, g32)], e64=[rel_max(a, b) for a, b in zip(ours, g64)],
                   ref=[rel_max(a, b) for a, b in zip(g32, g64)],
                   rows32=[rel_rows(a, b, s) for a, b, s in zip(ours, g32, shapes)],
                   rows_ref=[rel_rows(a, b, s) for a, b, s in zip(g32, g64, shapes)])
        if with_twin:
            twin, unpermute = permuted_twin(cpu_model, seed)
            g32t = unpermute(torch_grads(torch.float32, source=twin))
            out["twin32"] = [rel_max(a, b) for a, b in zip(g32t, g32)]
        if with_masks:
            g64m = torch_grads(torch.float64, masks)
            out["e64m"] = [rel_max(a, b) for a, b in zip(ours, g64m)]
        return out
    finally:
        lib.ocppo_set_gemm_mode(prev)
        torch.cuda.empty_cache()
