"""The GEMM family (fp32 FMA and tcgen05 3xTF32 engines) against torch fp64 on device."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))

CASES = [
    (0, 128, 256, 32, 1), (0, 256, 512, 256, 1), (0, 4096, 512, 256, 1), (0, 1000, 256, 96, 1),
    (0, 2048, 512, 2048, 1),
    (1, 512, 256, 512, 1), (1, 4096, 1024, 512, 1), (1, 640, 2048, 512, 1),
    (2, 256, 256, 4096, 4), (2, 512, 1024, 8192, 8), (2, 256, 512, 1000, 2), (2, 1024, 512, 16384, 9),
    # more tiles than SMs (persistent tensor-core kernel walks several tiles per CTA), ragged M
    (0, 32768, 1024, 512, 1), (1, 19999, 2048, 512, 1), (2, 512, 2048, 32768, 19),
]


@pytest.mark.parametrize("engine", [0, 1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("op,M,N,K,S", CASES)
def test_gemm_engine_vs_fp64(built_lib, engine, op, M, N, K, S):
    """Engines 6 / 7 are the fp16-pair storage of the learner: operands split into (hi, lo) fp16 tensors before the GEMM
    (conversion-free kernel), 7 also writes the output as a pair scaled by its a-priori bound and merges it back."""
    import gemm_diag
    if engine == 7 and op == 2:
        pytest.skip("weight gradients always leave in fp32")
    msg = gemm_diag.run(op, engine, M, N, K, S, "randn", seed=op * 7 + M)
    assert "BAD" not in msg and "rc=" not in msg, msg
    rel = float(msg.split("rel_err=")[1].split()[0])
    assert rel < 2e-5, msg            # fp32-grade: both engines, far inside the 1e-4 bar
    if op == 2:
        assert float(msg.split("bias_rel=")[1].split()[0]) < 2e-5, msg


# widths that are a multiple of 128 but not of 256 (the reference class's own default encoder_dims=(128, 64); 384-wide
# layers): the conversion-free kernel's 128-column tile.  Engines 1-5 have no such tile (those shapes run on the FMA kernel).
CASES_128 = [
    (0, 4096, 128, 256, 1), (0, 1000, 384, 128, 1), (0, 32768, 640, 512, 1),
    (1, 4096, 128, 512, 1), (1, 640, 384, 256, 1), (1, 19999, 128, 384, 1),
    (2, 256, 128, 4096, 4), (2, 128, 384, 8192, 8), (2, 384, 128, 16384, 9), (2, 640, 1152, 32768, 19),
]


@pytest.mark.parametrize("engine", [6, 7, 8])
@pytest.mark.parametrize("op,M,N,K,S", CASES_128)
def test_gemm_128_column_tile_vs_fp64(built_lib, engine, op, M, N, K, S):
    import gemm_diag
    if engine == 7 and op == 2:
        pytest.skip("weight gradients always leave in fp32")
    msg = gemm_diag.run(op, engine, M, N, K, S, "randn", seed=op * 7 + M)
    assert "BAD" not in msg and "rc=" not in msg, msg
    assert float(msg.split("rel_err=")[1].split()[0]) < 2e-5, msg
    if op == 2:
        assert float(msg.split("bias_rel=")[1].split()[0]) < 2e-5, msg


@pytest.mark.parametrize("kind", ["wide", "big", "small", "zero", "relu", "tiny+relu"])
@pytest.mark.parametrize("op,M,N,K,S", [(0, 1000, 512, 256, 1), (1, 640, 1024, 512, 1), (2, 256, 512, 8192, 4)])
@pytest.mark.parametrize("engine", [5, 6, 7, 8])
def test_gemm_fp16_engine_operand_ranges(built_lib, engine, op, M, N, K, S, kind):
    """Engine 5 scales every operand tensor by a power of two derived from its max-abs: magnitudes far from 1, nine
    decades of spread inside one tensor, and all-zero operands must neither overflow nor lose the result."""
    import gemm_diag
    if "+" in kind and op != 2:
        pytest.skip("mixed operand kinds are a WGRAD case")
    if engine == 7 and op == 2:
        pytest.skip("weight gradients always leave in fp32")
    msg = gemm_diag.run(op, engine, M, N, K, S, kind, seed=11 + op)
    assert "BAD" not in msg and "rc=" not in msg, msg
    assert float(msg.split("rel_err=")[1].split()[0]) < 2e-5, msg


def test_gemm_mode_switch_roundtrip(built_lib):
    default = built_lib.ocppo_get_gemm_mode()
    assert built_lib.ocppo_set_gemm_mode(1) == 0 and built_lib.ocppo_get_gemm_mode() == 1
    assert built_lib.ocppo_set_gemm_mode(0) == 0 and built_lib.ocppo_get_gemm_mode() == 0
    assert built_lib.ocppo_set_gemm_mode(2) == 0 and built_lib.ocppo_get_gemm_mode() == 2
    assert built_lib.ocppo_set_gemm_mode(3) == 0 and built_lib.ocppo_get_gemm_mode() == 3
    assert built_lib.ocppo_set_gemm_mode(4) == 0 and built_lib.ocppo_get_gemm_mode() == 4
    assert built_lib.ocppo_set_gemm_mode(5) == 0 and built_lib.ocppo_get_gemm_mode() == 5
    assert built_lib.ocppo_set_gemm_mode(6) == 0 and built_lib.ocppo_get_gemm_mode() == 6
    assert built_lib.ocppo_set_gemm_mode(default) == 0
    assert built_lib.ocppo_set_gemm_mode(7) == -1
