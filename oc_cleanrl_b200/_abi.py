"""ctypes binding of ``csrc/libocppo.so`` (the C ABI declared in ``include/ocppo.h``).

There is no fallback: if the shared library is missing or a call fails, this
module raises.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libocppo.so")

OCPPO_MAX_LAYERS = 8
OCPPO_MAX_ACTIONS = 31
OCPPO_NUM_STATS = 8
STAT_NAMES = ("pg_loss", "v_loss", "entropy", "loss", "old_approx_kl", "approx_kl", "clipfrac")


class OcppoError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [
        ("frames", C.c_int32),
        ("obs_dim", C.c_int32),
        ("n_actions", C.c_int32),
        ("n_encoder", C.c_int32),
        ("encoder_dims", C.c_int32 * OCPPO_MAX_LAYERS),
        ("n_decoder", C.c_int32),
        ("decoder_dims", C.c_int32 * OCPPO_MAX_LAYERS),
        ("device", C.c_int32),
        ("reserved", C.c_int32),
    ]


class HParams(C.Structure):
    _fields_ = [
        ("clip_coef", C.c_double),
        ("ent_coef", C.c_double),
        ("vf_coef", C.c_double),
        ("norm_adv", C.c_int32),
        ("clip_vloss", C.c_int32),
    ]


_P = C.c_void_p
_I64 = C.c_int64
_U64 = C.c_uint64
_I32 = C.c_int32
_F = C.c_float
_D = C.c_double

_SIGNATURES = {
    "ocppo_last_error": (C.c_char_p, []),
    "ocppo_version": (C.c_int, []),
    "ocppo_bringup_build": (C.c_int, []),
    "ocppo_launch_count": (_I64, []),
    "ocppo_launch_counts": (C.c_int, [C.POINTER(_I64), C.POINTER(_I64), _I32]),
    "ocppo_profile_enable": (C.c_int, [_I32]),
    "ocppo_profile_read": (C.c_int, [C.POINTER(_D), C.POINTER(_I64), _I32]),
    "ocppo_measure_fma_peak": (C.c_int, [C.POINTER(_D), _P]),
    "ocppo_set_gemm_mode": (C.c_int, [_I32]),
    "ocppo_get_gemm_mode": (C.c_int, []),
    "ocppo_gemm": (C.c_int, [_I32, _I32, _I32, _I32, _I64, _P, _I64, _P, _I64, _P, _I64, _P, _P, _P, _I32, _I64, _P]),
    "ocppo_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "ocppo_destroy": (C.c_int, [_P]),
    "ocppo_num_tensors": (C.c_int, [_P, C.POINTER(_I32), C.POINTER(_I64), C.POINTER(_I64)]),
    "ocppo_tensor_layout": (C.c_int, [_P, _I32, C.POINTER(_I64), C.POINTER(_I32), C.POINTER(_I32)]),
    "ocppo_bind_arena": (C.c_int, [_P, _P, _P, _P, _P]),
    "ocppo_workspace_bytes": (C.c_int, [_P, _I64, _I32, C.POINTER(_I64)]),
    "ocppo_bind_workspace": (C.c_int, [_P, _P, _I64, _I64, _I32]),
    "ocppo_policy_step": (C.c_int, [_P, _P, _I64, _U64, _U64, _P, _P, _P, _P, _P, _P]),
    "ocppo_policy_step_dev": (C.c_int, [_P, _P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "ocppo_value": (C.c_int, [_P, _P, _I64, _P, _P]),
    "ocppo_weights_version": (C.c_int, [_P, _I64]),
    "ocppo_weights_prepared": (C.c_int, [_P]),
    "ocppo_host_state_bytes": (_I64, []),
    "ocppo_host_state_save": (C.c_int, [_P, _P, _I64]),
    "ocppo_host_state_restore": (C.c_int, [_P, _P, _I64]),
    "ocppo_categorical": (C.c_int, [_P, _I64, _I32, _U64, _U64, _P, _P, _P, _P, _P]),
    "ocppo_philox_increment": (_U64, [_I64, _I32, _I32]),
    "ocppo_exponential": (C.c_int, [_P, _I64, _U64, _U64, _P]),
    "ocppo_gae": (C.c_int, [_P, _P, _P, _P, _P, _I64, _I64, _D, _D, _P, _P, _P]),
    "ocppo_gae_overlapped": (C.c_int, [_P, _P, _P, _P, _P, _I64, _I64, _D, _D, _P, _P, _P]),
    "ocppo_set_gae_variant": (C.c_int, [_I32]),
    "ocppo_minibatch_fwd_bwd": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _I64, C.POINTER(HParams), _P, _P]),
    "ocppo_clip_adam": (C.c_int, [_P, _F, _F, _D, _D, _D, _D, _I64, _P, _P]),
    "ocppo_adam_scalars_set": (C.c_int, [_P, _F, _F, _D, _D, _D, _D, _I64, _P]),
    "ocppo_philox_set": (C.c_int, [_P, _U64, _U64, _P]),
    "ocppo_clip_adam_dev": (C.c_int, [_P, _P, _P, _P]),
    "ocppo_trunk_forward": (C.c_int, [_P, _P, _I64, _P, _P]),
    "ocppo_explained_variance": (C.c_int, [_P, _P, _P, _I64, _P, _P]),
    "ocppo_explained_variance_static": (C.c_int, [_P, _P, _I64, _P, _P]),
    "ocppo_read_activation": (C.c_int, [_P, C.c_int32, _I64, _P, _P]),
    "ocppo_trunk_backward": (C.c_int, [_P, _P, _I64, _P, _P]),
    "ocppo_feed_create": (C.c_int, [_I32, _I32, _I32, _D, _D, _D, C.POINTER(_P)]),
    "ocppo_feed_destroy": (C.c_int, [_P]),
    "ocppo_feed_push": (C.c_int, [_P, _P, _P, _P, _P]),
    "ocppo_feed_normalize_reward": (C.c_int, [_P, _P, _P, _P, _P]),
    "ocppo_feed_state": (C.c_int, [_P, _P, _P]),
}

_lib = None


def lib():
    """Load libocppo.so once; raise loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OcppoError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  oc_cleanrl_b200 has no CPU or eager-torch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


KERNEL_CLASSES = ("gae", "gemm_fwd", "gemm_dgrad", "gemm_wgrad", "heads", "aux", "clip_adam", "gemm_cuda_core")


def profile_read():
    """-> ({class: ms}, {class: timed launches}) since the last read."""
    ms = (_D * 8)()
    cnt = (_I64 * 8)()
    check(lib().ocppo_profile_read(ms, cnt, 8))
    return ({k: ms[i] for i, k in enumerate(KERNEL_CLASSES)}, {k: cnt[i] for i, k in enumerate(KERNEL_CLASSES)})


def launch_counts():
    """-> ({class: kernels launched since load}, {class: launched under graph capture while profiling, hence untimed})."""
    n = (_I64 * 8)()
    u = (_I64 * 8)()
    check(lib().ocppo_launch_counts(n, u, 8))
    return ({k: n[i] for i, k in enumerate(KERNEL_CLASSES)}, {k: u[i] for i, k in enumerate(KERNEL_CLASSES)})


def exported_symbols():
    return tuple(_SIGNATURES)


def check(status: int):
    if status != 0:
        msg = lib().ocppo_last_error()
        raise OcppoError(f"libocppo error {status}: {msg.decode() if msg else '?'}")


def ptr(t):
    """Device (or null) pointer of a torch tensor."""
    return None if t is None else C.c_void_p(t.data_ptr())


def make_config(frames, obs_dim, n_actions, encoder_dims, decoder_dims, device):
    if len(encoder_dims) > OCPPO_MAX_LAYERS or len(decoder_dims) > OCPPO_MAX_LAYERS:
        raise OcppoError(f"at most {OCPPO_MAX_LAYERS} encoder and decoder layers are supported")
    cfg = Config()
    cfg.frames, cfg.obs_dim, cfg.n_actions = int(frames), int(obs_dim), int(n_actions)
    cfg.n_encoder, cfg.n_decoder = len(encoder_dims), len(decoder_dims)
    for i, d in enumerate(encoder_dims):
        cfg.encoder_dims[i] = int(d)
    for i, d in enumerate(decoder_dims):
        cfg.decoder_dims[i] = int(d)
    cfg.device = int(device)
    return cfg


class CapturedCall:
    """A CUDA graph of captured libocppo calls plus the handle's host-side bookkeeping as the captured calls left it
    (``ocppo_host_state_save``): ``replay()`` launches the graph and restores the bookkeeping."""

    def __init__(self, handle_raw, graph):
        self.handle_raw, self.graph = handle_raw, graph
        n = lib().ocppo_host_state_bytes()
        self.state = C.create_string_buffer(n)
        check(lib().ocppo_host_state_save(handle_raw, self.state, n))

    def replay(self):
        self.graph.replay()
        check(lib().ocppo_host_state_restore(self.handle_raw, self.state, len(self.state)))


def capture(handle_raw, device, fn):
    """Run ``fn(stream_ptr)`` (libocppo calls on the given stream) under CUDA-graph capture on a side stream of `device`
    and return the :class:`CapturedCall`.  The graph has not run yet: call ``replay()``."""
    import torch
    cur = torch.cuda.current_stream(device)
    side = torch.cuda.Stream(device)
    side.wait_stream(cur)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
            fn(C.c_void_p(side.cuda_stream))
    cur.wait_stream(side)
    return CapturedCall(handle_raw, g)


class Handle:
    """RAII wrapper around ``ocppo_handle*``."""

    def __init__(self, cfg: Config):
        self._h = _P()
        check(lib().ocppo_create(C.byref(cfg), C.byref(self._h)))
        self.cfg = cfg
        n, arena, params = _I32(), _I64(), _I64()
        check(lib().ocppo_num_tensors(self._h, C.byref(n), C.byref(arena), C.byref(params)))
        self.n_tensors, self.arena_floats, self.param_floats = n.value, arena.value, params.value
        self.layout = []
        for i in range(self.n_tensors):
            off, rows, cols = _I64(), _I32(), _I32()
            check(lib().ocppo_tensor_layout(self._h, i, C.byref(off), C.byref(rows), C.byref(cols)))
            self.layout.append((off.value, rows.value, cols.value))

    @property
    def raw(self):
        return self._h

    def workspace_bytes(self, max_rows: int, training: bool) -> int:
        out = _I64()
        check(lib().ocppo_workspace_bytes(self._h, int(max_rows), int(training), C.byref(out)))
        return out.value

    def close(self):
        if self._h:
            lib().ocppo_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
