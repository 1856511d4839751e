"""ORACLE (test infrastructure, never on the product path): dropout-mask generator.

The reference draws dropout masks from Paddle's device RNG inside
`get_rng_state_tracker().rng_state(...)` (hybrid_model.py:280, 546, 556; seeds
ppfleetx/utils/env.py:27-46). Paddle is not installable here, so its stream cannot be matched
bit for bit; the library therefore DEFINES the mask as a counter-based function
(fleetx_b200/csrc/philox.cuh) and this file restates that definition in numpy so the CUDA
masks can be compared bit for bit. Parity of the mask with Paddle is unpinned by construction;
what is pinned is (seed, offset, index) -> mask between CUDA and this restatement.

  key     = (seed lo32, seed hi32)
  counter = (group lo32, group hi32, offset lo32, offset hi32)
  Philox4x32-10 -> 4 words -> 8 x 16-bit lanes (element 2i = low half of word i, 2i+1 = high)
  keep    <=>  (lane & 0x7fff) >= round(p * 32768)
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(seed: int, offset: int, groups: np.ndarray) -> np.ndarray:
    """groups: uint64 array [...]. Returns uint32 array [..., 4]."""
    g = np.asarray(groups, dtype=np.uint64)
    c0 = g & _MASK32
    c1 = g >> np.uint64(32)
    c2 = np.full_like(g, np.uint64(offset & 0xFFFFFFFF))
    c3 = np.full_like(g, np.uint64((offset >> 32) & 0xFFFFFFFF))
    k0 = seed & 0xFFFFFFFF
    k1 = (seed >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0  # both factors < 2^32, so the product fits in uint64
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK32
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, lo1, n2, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def dropout_threshold(p: float) -> int:
    t = np.float32(p) * np.float32(32768.0) + np.float32(0.5)
    return int(min(max(float(t), 0.0), 32767.0))


def _keep_from_words(words: np.ndarray, p: float) -> np.ndarray:
    """words uint32 [..., 4] -> keep uint8 [..., 8]."""
    lo = (words & np.uint32(0x7FFF)).astype(np.uint32)
    hi = ((words >> np.uint32(16)) & np.uint32(0x7FFF)).astype(np.uint32)
    draws = np.stack([lo, hi], axis=-1).reshape(*words.shape[:-1], 8)
    return (draws >= np.uint32(dropout_threshold(p))).astype(np.uint8)


def dropout_mask(seed: int, offset: int, n_elements: int, p: float) -> np.ndarray:
    """Mask for a flat tensor of n_elements (the [rows, h] activations of dropout1/dropout2)."""
    if p <= 0.0:
        return np.ones(n_elements, dtype=np.uint8)
    ngroups = (n_elements + 7) // 8
    words = philox4x32_10(seed, offset, np.arange(ngroups, dtype=np.uint64))
    return _keep_from_words(words, p).reshape(-1)[:n_elements]


def attn_dropout_mask(seed: int, offset: int, b: int, nh: int, s: int, p: float) -> np.ndarray:
    """Mask for attention probabilities, uint8 [b, nh, s, s].

    group index = ((batch * nh + head) * s + row) * ceil(s / 8) + col // 8 (attention.cu).
    """
    if p <= 0.0:
        return np.ones((b, nh, s, s), dtype=np.uint8)
    s8 = (s + 7) // 8
    rows = np.arange(b * nh * s, dtype=np.uint64)[:, None] * np.uint64(s8)
    groups = rows + np.arange(s8, dtype=np.uint64)[None, :]
    words = philox4x32_10(seed, offset, groups)  # [b*nh*s, s8, 4]
    keep = _keep_from_words(words, p).reshape(b * nh * s, s8 * 8)[:, :s]
    return keep.reshape(b, nh, s, s)
