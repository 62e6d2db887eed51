"""Numpy restatement of the in-kernel counter-based RNG contract.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  This is not reference
arithmetic -- the reference draws eps with torch's own generator
(linear_variational.py:164) -- it is the specification of the stream the CUDA
kernels generate in registers (csrc/bt_rng.cuh), restated so tests can
predict it:

  Philox4x32-10, key = (seed_lo, seed_hi),
  counter = (block_lo, block_hi, sample_id, stream_id)
  normals : block = elem_idx // 8; word r_i -> (z_2i, z_2i+1) by Box-Muller with
            u0 = ((r_i >> 16) + 0.5) * 2^-16 (radius), u1 = ((r_i & 0xFFFF) + 0.5) * 2^-16 (angle):
            z_2i = sqrt(-2 ln u0) cos(2 pi (u1 - 1/2)), z_2i+1 = sqrt(-2 ln u0) sin(2 pi (u1 - 1/2))
  signs   : block = elem_idx // 128; word r_((elem_idx // 32) % 4); element j = elem_idx % 32 reads bit
            (j // 2) + 16 * (j % 2) of that word, 1 -> -1 (the two halves of a packed bf16 pair sit 16 bits
            apart, so one shift + mask flips both).  elem_idx of an activation element (b, c, spatial...)
            is its CHANNELS-LAST position ((b, spatial...), c), whatever the tensor's memory layout
            (sign_like() below).
  stream_id = layer_id << 4 | kind, kind: 0 eps_weight, 1 eps_bias, 2 sign_in, 3 sign_out
"""
from __future__ import annotations

import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)

KIND_EPS_W, KIND_EPS_B, KIND_SIGN_IN, KIND_SIGN_OUT = 0, 1, 2, 3


def stream_id(layer_id: int, kind: int) -> int:
    return ((layer_id << 4) | kind) & 0xFFFFFFFF


def philox4x32_10(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32-10; c* are uint32 arrays (broadcastable)."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & MASK for c in np.broadcast_arrays(c0, c1, c2, c3))
    k0 &= 0xFFFFFFFF
    k1 &= 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)) & MASK, lo1, (hi0 ^ c3 ^ np.uint64(k1)) & MASK, lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def _split_seed(seed: int):
    return seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF


def normal(n: int, seed: int, sample_id: int, stream: int, start: int = 0) -> np.ndarray:
    """eps[start:start+n] of the (seed, sample_id, stream) normal stream, float64."""
    first, last = (start // 8) if n else 0, (start + n + 7) // 8
    blocks = np.arange(first, last, dtype=np.uint64)
    k0, k1 = _split_seed(seed)
    r = philox4x32_10(blocks & MASK, blocks >> np.uint64(32), np.uint64(sample_id & 0xFFFFFFFF),
                      np.uint64(stream & 0xFFFFFFFF), k0, k1)
    z = np.empty((len(blocks), 8), dtype=np.float64)
    for j, w in enumerate(r):
        u0 = ((w >> np.uint32(16)).astype(np.float64) + 0.5) * 2.0 ** -16
        u1 = ((w & np.uint32(0xFFFF)).astype(np.float64) + 0.5) * 2.0 ** -16
        rad = np.sqrt(-2.0 * np.log(u0))
        ang = 2.0 * np.pi * (u1 - 0.5)
        z[:, 2 * j] = rad * np.cos(ang)
        z[:, 2 * j + 1] = rad * np.sin(ang)
    z = z.reshape(-1)
    off = start - first * 8
    return z[off:off + n]


def sign(n: int, seed: int, sample_id: int, stream: int, start: int = 0) -> np.ndarray:
    """Rademacher +-1 for elements [start, start+n) of the sign stream, float32."""
    idx = np.arange(start, start + n, dtype=np.uint64)
    blocks = idx // np.uint64(128)
    ub, inv = np.unique(blocks, return_inverse=True)
    k0, k1 = _split_seed(seed)
    r = philox4x32_10(ub & MASK, ub >> np.uint64(32), np.uint64(sample_id & 0xFFFFFFFF),
                      np.uint64(stream & 0xFFFFFFFF), k0, k1)
    words = np.stack(r, axis=1)  # [nblocks, 4]
    within = (idx % np.uint64(128)).astype(np.int64)
    w = words[inv, within // 32]
    j = within % 32
    bit = (w >> ((j // 2) + 16 * (j % 2)).astype(np.uint32)) & np.uint32(1)
    return np.where(bit == 1, -1.0, 1.0).astype(np.float32)


def sign_like(shape, seed: int, sample_id: int, stream: int) -> np.ndarray:
    """+-1 tensor of logical shape [B, C, *spatial] holding the signs the kernels use for each element:
    the stream runs over the tensor in channels-last order ((b, spatial...), c)."""
    shape = tuple(int(v) for v in shape)
    n = int(np.prod(shape))
    flat = sign(n, seed, sample_id, stream)
    if len(shape) <= 2:
        return flat.reshape(shape)
    cl = flat.reshape((shape[0],) + shape[2:] + (shape[1],))
    return np.ascontiguousarray(np.moveaxis(cl, -1, 1))
