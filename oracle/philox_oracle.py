"""CPU restatement of the in-kernel noise generator.  TEST INFRASTRUCTURE ONLY.

The reference draws its source noise with ``torch.randn_like`` (hifigan.py:223); the CUDA path can instead draw
standard normals inside the source kernel ("philox" noise, csrc/aux_kernels.cu: philox_normal).  That generator
has no counterpart in the reference, so its oracle is the published algorithm:

  * Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11; Random123
    library).  Pinned by the Random123 known-answer vectors (tests/test_oracle.py).
  * Box-Muller on the two word pairs of a block: u1 = (w_a + 1) * 2^-32 in (0, 1], u2 = w_b * 2^-32,
    z_even = sqrt(-2 ln u1) cos(2 pi u2), z_odd = sqrt(-2 ln u1) sin(2 pi u2).

Draw ``s`` of item ``b`` of a call with key ``seed`` is element ``(b*n + s) & 3`` of block ``(b*n + s) >> 2``
(counter = (block & 0xffffffff, block >> 32, 0, 0), key = (seed & 0xffffffff, seed >> 32)), with b the GLOBAL item
index: the waveform therefore does not depend on how a batch is cut into passes.
"""
from __future__ import annotations

import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = np.uint32(0x9E3779B9)
W1 = np.uint32(0xBB67AE85)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32 with 10 rounds.  All arguments broadcastable uint32 arrays; returns 4 uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint32) for v in (c0, c1, c2, c3))
    k0 = np.asarray(k0, dtype=np.uint32)
    k1 = np.asarray(k1, dtype=np.uint32)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = (k0 + W0).astype(np.uint32)
            k1 = (k1 + W1).astype(np.uint32)
    return c0, c1, c2, c3


def philox_normal(seed: int, idx) -> np.ndarray:
    """Standard-normal draw number ``idx`` (uint64 array) of key ``seed`` - what philox_normal() in
    csrc/aux_kernels.cu returns up to the device's fast-math log / sincospi rounding (~1e-6 relative)."""
    idx = np.asarray(idx, dtype=np.uint64)
    blk = idx >> np.uint64(2)
    r = philox4x32_10((blk & np.uint64(0xFFFFFFFF)).astype(np.uint32), (blk >> np.uint64(32)).astype(np.uint32), 0, 0,
                      np.uint32(seed & 0xFFFFFFFF), np.uint32((seed >> 32) & 0xFFFFFFFF))
    lane = (idx & np.uint64(3)).astype(np.int64)
    ua = np.where(lane < 2, r[0], r[2]).astype(np.float64)
    ub = np.where(lane < 2, r[1], r[3]).astype(np.float64)
    # the kernel forms u1, u2 in fp32: (float)ua rounds to 24 bits first
    u1 = (ua.astype(np.float32) + np.float32(1.0)) * np.float32(2.3283064365386963e-10)
    u2 = ub.astype(np.float32) * np.float32(2.3283064365386963e-10)
    rad = np.sqrt(-2.0 * np.log(u1.astype(np.float64)))
    ang = 2.0 * np.pi * u2.astype(np.float64)
    return np.where(lane & 1, rad * np.sin(ang), rad * np.cos(ang))


def source_noise(seed: int, item0: int, B: int, n: int) -> np.ndarray:
    """(B, n) draws of items item0 .. item0+B-1 with n samples each."""
    idx = (np.arange(B, dtype=np.uint64)[:, None] + np.uint64(item0)) * np.uint64(n) + np.arange(n, dtype=np.uint64)[None, :]
    return philox_normal(seed, idx)


# Random123 known-answer vectors (kat_vectors, "philox4x32 10"): (counter, key) -> output
KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff),
     (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]
