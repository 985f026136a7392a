"""``ocppo_explained_variance_static`` -- the allocation-free variant of ``ocppo_explained_variance`` (scratch in a file-scope
``static __device__`` array inside the library; no counterpart in the reference)
-- against the existing entry point, through the C ABI, on two calls in a row.  Collected last on purpose
(``zz``).  Needs a B200: ``-m gpu``."""
import ctypes as C

import numpy as np
import pytest
import torch

from oc_cleanrl_b200.synthetic import stub_envs

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4          # the bar tests/test_gpu_rollout.py holds ocppo_explained_variance to


@pytest.fixture(scope="module")
def dev(built_lib):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _existing(agent, v, r):
    """ocppo_explained_variance: needs a handle with a bound training workspace (its scratch)."""
    from oc_cleanrl_b200 import _abi
    agent._ensure_workspace(1, training=True)
    out = torch.empty(2, dtype=torch.float32, device=v.device)
    _abi.check(_abi.lib().ocppo_explained_variance(agent._handle.raw, _abi.ptr(v), _abi.ptr(r), v.numel(), _abi.ptr(out),
                                                   agent._stream()))
    return out


def _static(v, r, out):
    from oc_cleanrl_b200 import _abi
    stream = C.c_void_p(torch.cuda.current_stream(v.device).cuda_stream)
    return _abi.lib().ocppo_explained_variance_static(_abi.ptr(v), _abi.ptr(r), v.numel(), _abi.ptr(out), stream)


def test_static_scratch_variant_equals_the_workspace_variant_on_two_calls_in_a_row(dev, built_lib):
    from oc_cleanrl_b200.architectures.ppo import PPObj
    torch.manual_seed(1)
    agent = PPObj(stub_envs(4, 6, 6), dev, (32, 64), (32,)).to(dev)
    rs = np.random.RandomState(0)
    # the second call is SMALLER than the first: fewer blocks, so stale partial sums of the first call sit in the array
    # beyond what the second one writes -- and must not be read
    cases = [(128 * 4096, 3.0), (1000, -2.0)]
    inputs = [(torch.as_tensor(rs.randn(n).astype(np.float32) + off).to(dev),
               torch.as_tensor((rs.randn(n) * 2.0).astype(np.float32)).to(dev)) for n, off in cases]
    want = [_existing(agent, v, r).cpu().numpy().astype(np.float64) for v, r in inputs]
    outs = [torch.full((2,), float("nan"), device=dev) for _ in cases]      # outputs allocated BEFORE the calls under test
    torch.cuda.synchronize()
    for (v, r), out in zip(inputs, outs):                                   # two calls in a row, same stream, no sync between
        assert _static(v, r, out) == 0, built_lib.ocppo_last_error()
    torch.cuda.synchronize()
    for (v, r), out, w in zip(inputs, outs, want):
        got = out.cpu().numpy().astype(np.float64)
        assert np.all(np.isfinite(got))
        assert np.allclose(got, w, rtol=REL_TOL, atol=0.0), (got, w)
        # and against the script's own line (ppo_atari_oc.py:620-623: np.var on host copies)
        y_pred, y_true = v.cpu().numpy(), r.cpu().numpy()
        assert np.allclose(got, [np.var(y_true), np.var(y_true - y_pred)], rtol=REL_TOL)
    # a third call repeating the first must reproduce it bit for bit (fixed summation order, shared array overwritten)
    again = torch.empty(2, device=dev)
    assert _static(*inputs[0], again) == 0
    torch.cuda.synchronize()
    assert torch.equal(again, outs[0])


def test_static_scratch_variant_rejects_sizes_beyond_its_maximum(dev, built_lib):
    v = torch.zeros(16, device=dev)
    out = torch.empty(2, device=dev)
    from oc_cleanrl_b200 import _abi
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    too_many = (1 << 28) + 1                      # OCPPO_EV_STATIC_MAX_N + 1: rejected before anything is launched or read
    rc = built_lib.ocppo_explained_variance_static(_abi.ptr(v), _abi.ptr(v), too_many, _abi.ptr(out), stream)
    assert rc == -1
    msg = built_lib.ocppo_last_error().decode()
    assert "maximum" in msg and str(1 << 28) in msg, msg
    assert built_lib.ocppo_explained_variance_static(_abi.ptr(v), _abi.ptr(v), 0, _abi.ptr(out), stream) == -1
    assert built_lib.ocppo_explained_variance_static(None, _abi.ptr(v), 16, _abi.ptr(out), stream) == -1
    cpu = torch.zeros(16)
    assert built_lib.ocppo_explained_variance_static(C.c_void_p(cpu.data_ptr()), _abi.ptr(v), 16, _abi.ptr(out), stream) == -1
    torch.cuda.synchronize()
