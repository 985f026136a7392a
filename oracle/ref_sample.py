"""CPU restatement of torch-CUDA ``Categorical(logits).sample()`` for one draw.

TEST INFRASTRUCTURE -- see ``oracle/__init__.py``.

Reference call site: ``cleanrl/architectures/ppo.py:92-94`` (``probs.sample()``).
The arithmetic lives in the reference's pinned third-party dependency, PyTorch
(``torch==1.12.1`` in requirements/requirements.txt:68; 2.11.0+cu128 here):

* ``torch/distributions/categorical.py:78``      logits -= logsumexp(logits)
* ``torch/distributions/categorical.py:144-149`` multinomial(probs, 1, True)
* ATen ``multinomial`` single-draw path          argmax(probs / q), q ~ Exp(1)
* ``ATen/native/cuda/DistributionTemplates.h:50-90,428-441``
      exponential_ on CUDA = Philox4x32-10 ``curand_uniform4`` per thread,
      thread ``idx`` = Philox subsequence, grid-stride element mapping
* ``ATen/core/TransformationHelper.h:129-146``   q = -log(u), u>=1-eps/2 -> eps/2
* Philox4x32-10 itself: Salmon et al., "Parallel random numbers: as easy as
  1, 2, 3" (SC'11); constants as ``ATen/core/PhiloxRNGEngine.h:229-232``.
"""
from __future__ import annotations

import numpy as np

PHILOX_M0 = np.uint64(0xD2511F53)
PHILOX_M1 = np.uint64(0xCD9E8D57)
PHILOX_W0 = 0x9E3779B9
PHILOX_W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)

TORCH_BLOCK = 256          # block_size_bound, DistributionTemplates.h:36
TORCH_UNROLL = 4           # float4 per curand_uniform4 call


def philox4x32_10(counter, key):
    """Vectorised Philox4x32-10.  counter: (...,4) uint32, key: (...,2) uint32."""
    c = [np.asarray(counter[..., i], dtype=np.uint64) for i in range(4)]
    k0 = np.asarray(key[..., 0], dtype=np.uint64)
    k1 = np.asarray(key[..., 1], dtype=np.uint64)
    for rnd in range(10):
        p0 = PHILOX_M0 * c[0]
        p1 = PHILOX_M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        c = [(hi1 ^ c[1] ^ k0) & MASK32, lo1, (hi0 ^ c[3] ^ k1) & MASK32, lo0]
        if rnd != 9:
            k0 = (k0 + np.uint64(PHILOX_W0)) & MASK32
            k1 = (k1 + np.uint64(PHILOX_W1)) & MASK32
    return np.stack(c, axis=-1).astype(np.uint32)


def torch_cuda_grid(numel, sm_count=148, max_threads_per_sm=2048):
    """calc_execution_policy, DistributionTemplates.h:50-62.
    Returns (grid, counter_offset)."""
    grid = (numel + TORCH_BLOCK - 1) // TORCH_BLOCK
    grid = min(sm_count * (max_threads_per_sm // TORCH_BLOCK), grid)
    counter_offset = ((numel - 1) // (TORCH_BLOCK * grid * TORCH_UNROLL) + 1) * 4
    return grid, counter_offset


def torch_cuda_uniform_bits(numel, seed, offset, sm_count=148, max_threads_per_sm=2048):
    """uint32 Philox word that torch's CUDA ``uniform_and_transform`` feeds to
    element ``li`` (DistributionTemplates.h:65-90): thread idx = li % G,
    lane ii = (li // G) % 4, curand4 call number n = li // (4 G), counter =
    (offset/4 + n, subsequence idx)."""
    grid, _ = torch_cuda_grid(numel, sm_count, max_threads_per_sm)
    G = grid * TORCH_BLOCK
    li = np.arange(numel, dtype=np.uint64)
    idx = li % np.uint64(G)
    ii = ((li // np.uint64(G)) % np.uint64(4)).astype(np.int64)
    call = li // np.uint64(4 * G)
    ctr64 = np.uint64(offset // 4) + call
    counter = np.stack([ctr64 & MASK32, ctr64 >> np.uint64(32), idx & MASK32, idx >> np.uint64(32)],
                       axis=-1).astype(np.uint32)
    key = np.empty((numel, 2), np.uint32)
    key[:, 0] = np.uint32(seed & 0xFFFFFFFF)
    key[:, 1] = np.uint32((seed >> 32) & 0xFFFFFFFF)
    words = philox4x32_10(counter, key)
    return words[np.arange(numel), ii]


def curand_uniform_from_bits(bits):
    """curand_uniform: x * 2^-32 + 2^-33, in (0, 1] (curand_uniform.h)."""
    x = bits.astype(np.float32)                      # cvt.rn.f32.u32
    return x * np.float32(2.3283064e-10) + np.float32(2.3283064e-10 / 2.0)


def torch_cuda_exponential(u):
    """TransformationHelper.h:129-146, lambda = 1, fp32."""
    eps = np.finfo(np.float32).eps
    edge = u >= np.float32(1.0) - np.float32(eps / 2)
    with np.errstate(divide="ignore"):
        lg = np.where(edge, np.float32(-eps / 2), np.log(u.astype(np.float32)))
    return (np.float32(-1.0) * lg).astype(np.float32)


def ref_categorical_sample(logits, seed, offset, sm_count=148, max_threads_per_sm=2048):
    """logits (B,A) float32 -> dict(action int64 (B,), logp, entropy, q, probs).

    categorical.py:78 (normalise), :144-149 (sample), :151-157 (log_prob),
    :159-163 (entropy)."""
    import torch
    z = torch.as_tensor(np.asarray(logits, dtype=np.float32))
    B, A = z.shape
    norm = z - z.logsumexp(dim=-1, keepdim=True)
    probs = torch.softmax(norm, dim=-1)
    bits = torch_cuda_uniform_bits(B * A, seed, offset, sm_count, max_threads_per_sm)
    q = torch_cuda_exponential(curand_uniform_from_bits(bits)).reshape(B, A)
    ratio = probs.numpy() / q
    action = np.argmax(ratio, axis=-1).astype(np.int64)       # first index on ties
    logp = norm.gather(-1, torch.as_tensor(action).unsqueeze(-1)).squeeze(-1)
    min_real = torch.finfo(norm.dtype).min
    entropy = -(torch.clamp(norm, min=min_real) * probs).sum(-1)
    return dict(action=action, logp=logp.numpy(), entropy=entropy.numpy(), q=q,
                probs=probs.numpy(), norm_logits=norm.numpy())


def philox_offset_increment(numel, sm_count=148, max_threads_per_sm=2048):
    """How far one ``exponential_`` call advances the generator offset."""
    return torch_cuda_grid(numel, sm_count, max_threads_per_sm)[1]
