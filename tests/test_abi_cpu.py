"""The C-ABI library loads and exports every symbol include/ocppo.h declares;
host-side (no-GPU) behaviour of the boundary: layout, sizes, error paths."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from oracle import ref_ppo, ref_sample


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ocppo.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ocppo_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported_and_bound(built_lib):
    from oc_cleanrl_b200 import _abi
    declared = _declared_symbols()
    assert len(declared) >= 18
    for name in declared:
        assert hasattr(built_lib, name), f"{name} declared in ocppo.h but not exported"
    assert sorted(_abi.exported_symbols()) == declared
    assert built_lib.ocppo_version() == 200


def test_library_is_sm100a_only(built_lib):
    import subprocess
    from oc_cleanrl_b200 import _abi
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", _abi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, archs


@pytest.mark.parametrize("F,D,A,enc,dec", [
    (4, 6, 6, (256, 512, 1024, 512), (512,)),      # Pong (config 2)
    (4, 88, 6, (256, 512, 1024, 512), (512,)),     # SpaceInvaders
    (4, 220, 4, (256, 512, 1024, 512), (512,)),    # Breakout
    (4, 6, 6, (128, 64), (32,)),                   # PPObj's own ctor defaults
    (2, 5, 3, (7,), (9, 4)),
])
def test_arena_layout_matches_parameters_order(built_lib, F, D, A, enc, dec):
    from oc_cleanrl_b200 import _abi
    h = _abi.Handle(_abi.make_config(F, D, A, enc, dec, device=-1))
    agent = ref_ppo.make_agent(F, D, A, enc, dec)
    params = list(agent.parameters())
    assert h.n_tensors == len(params)
    assert h.param_floats == sum(p.numel() for p in params)
    end = 0
    for p, (off, rows, cols) in zip(params, h.layout):
        assert rows * cols == p.numel() and rows == p.shape[0]
        assert off % 32 == 0 and off >= end
        end = off + rows * cols
    assert h.arena_floats >= end and h.arena_floats % 32 == 0
    if (F, D, A) == (4, 6, 6) and enc == (256, 512, 1024, 512):
        assert h.param_floats == 2236167
    assert h.workspace_bytes(256, True) > h.workspace_bytes(256, False) > 0


def test_error_paths_without_gpu(built_lib):
    from oc_cleanrl_b200 import _abi
    with pytest.raises(_abi.OcppoError, match="n_actions"):
        _abi.Handle(_abi.make_config(4, 6, 0, (8,), (8,), device=-1))
    with pytest.raises(_abi.OcppoError, match="trunk layer"):
        _abi.Handle(_abi.make_config(4, 6, 6, (), (), device=-1))
    h = _abi.Handle(_abi.make_config(4, 6, 6, (8,), (8,), device=-1))
    rc = built_lib.ocppo_value(h.raw, None, 1, None, None)
    assert rc == -4 and b"layout-only" in built_lib.ocppo_last_error()       # OCPPO_ESTATE, no CPU fallback
    rc = built_lib.ocppo_gae(None, None, None, None, None, -1, 4, 0.99, 0.95, None, None, None)
    assert rc == -1
    assert built_lib.ocppo_gae(None, None, None, None, None, 0, 0, 0.99, 0.95, None, None, None) == 0   # empty
    if not torch.cuda.is_available():
        with pytest.raises(_abi.OcppoError):
            _abi.Handle(_abi.make_config(4, 6, 6, (8,), (8,), device=0))      # needs a real sm_100 device


def test_philox_increment_matches_torch_policy(built_lib):
    for numel in [1, 6, 255, 256, 24576, 303104, 1212416, 1212417, 5_000_000]:
        assert built_lib.ocppo_philox_increment(numel, 148, 2048) == ref_sample.philox_offset_increment(numel)
    assert built_lib.ocppo_philox_increment(0, 148, 2048) == 0


def test_product_agent_refuses_cpu_compute(built_lib):
    from oc_cleanrl_b200 import _abi
    from oc_cleanrl_b200.architectures.ppo import PPObj
    torch.manual_seed(42)
    agent = PPObj(ref_ppo.stub_envs(4, 6, 6), "cpu", (16,), (8,))
    oracle = ref_ppo.make_agent(4, 6, 6, (16,), (8,), seed=42)
    assert list(agent.state_dict().keys()) == list(oracle.state_dict().keys())
    for a, b in zip(agent.parameters(), oracle.parameters()):
        assert torch.equal(a, b)                      # same init stream as the reference
    with pytest.raises(_abi.OcppoError, match="no CPU path"):
        agent.get_value(torch.zeros(2, 4, 6))
    assert agent.predict(np.zeros((3, 4, 6), np.float32))[0].shape == (3,)    # Predictor surface (eager, eval only)
