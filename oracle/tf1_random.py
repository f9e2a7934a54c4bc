"""CPU oracle for TF 1.x `tf.random_normal`: Philox4x32-10 + Box-Muller, as TensorFlow's CPU kernel evaluates it.

TEST INFRASTRUCTURE ONLY (see oracle/tf1_graph.py): nothing in the product path imports this module.

PARITY UNPINNED against TensorFlow itself: TF 1.x is not installable here (README.md:6-8 of the reference leaves it
unpinned; no wheel for this Python) and the reference ships no fixture with the values `distributed.py:85-86` draws.
What pins this file (tests/test_oracle.py):
  * the Philox4x32-10 block function against the three known-answer vectors published with Random123 (the generator
    TensorFlow's `PhiloxRandom` implements: same multipliers 0xD2511F53 / 0xCD9E8D57, Weyl constants 0x9E3779B9 /
    0xBB67AE85, ten rounds);
  * statistical checks of the normal stream (mean, variance, tail), determinism and the 4-per-block layout.

Restated from the published algorithm of the third-party dependency [TF1] (tensorflow/core/lib/random/
philox_random.h, random_distributions.h, kernels/random_op.cc); the reference's call sites are
  distributed/distributed.py:83      tf.set_random_seed(2)              -> graph seed = 2 ("seed" attr of the op)
  distributed/distributed.py:85-86   tf.random_normal([784,100]) / ([100,10]), no op seed
                                     -> "seed2" attr = the graph's op counter when the op was created [TF1]
  PhiloxRandom(seed, seed2): key = (lo32(seed), hi32(seed)); counter = (0, 0, lo32(seed2), hi32(seed2))
  output[4g .. 4g+3] = BoxMuller(philox(counter + g))   (each block of four uint32 makes four normals)
  Uint32ToFloat(x)   = bitcast(0x3F800000 | (x & 0x7FFFFF)) - 1.0          in [0, 1)
  BoxMullerFloat     : u1 = max(Uint32ToFloat(x0), 1e-7); v1 = 2*pi*Uint32ToFloat(x1); r = sqrt(-2 ln u1);
                       f0 = sin(v1) * r; f1 = cos(v1) * r
The op counter that becomes seed2 depends on the TensorFlow version (how many graph nodes the lines before :85 create),
which is why the drop-in takes the pair explicitly (`distributed.py --tf_op_seeds`).
"""
from __future__ import annotations

import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(counter, key):
    """counter: uint32 array [..., 4]; key: two uint32.  Returns uint32 [..., 4] (one Philox block per counter)."""
    c = np.asarray(counter, dtype=np.uint64).copy()
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[..., 0]
        p1 = M1 * c[..., 2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        n0 = hi1 ^ c[..., 1] ^ np.uint64(k0)
        n2 = hi0 ^ c[..., 3] ^ np.uint64(k1)
        c = np.stack([n0, lo1, n2, lo0], axis=-1)
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return c.astype(np.uint32)


def _uint32_to_float(x):
    bits = (np.asarray(x, dtype=np.uint32) & np.uint32(0x7FFFFF)) | np.uint32(0x3F800000)
    return bits.view(np.float32) - np.float32(1.0)


def random_normal(shape, seed, seed2, mean=0.0, stddev=1.0):
    """float32 array of `shape` with the values tf.random_normal(shape, mean, stddev) yields for op attrs (seed, seed2)."""
    n = int(np.prod(shape))
    groups = (n + 3) // 4
    g = np.arange(groups, dtype=np.uint64)
    ctr = np.zeros((groups, 4), dtype=np.uint64)
    ctr[:, 0] = g & MASK                                  # Skip(g): 64-bit add into counter[0..1]
    ctr[:, 1] = g >> np.uint64(32)
    ctr[:, 2] = np.uint64(int(seed2) & 0xFFFFFFFF)
    ctr[:, 3] = np.uint64((int(seed2) >> 32) & 0xFFFFFFFF)
    r = philox4x32_10(ctr, (int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF))
    out = np.empty((groups, 4), dtype=np.float32)
    for i in (0, 2):
        u1 = np.maximum(_uint32_to_float(r[:, i]), np.float32(1.0e-7))
        v1 = np.float32(2.0 * np.pi) * _uint32_to_float(r[:, i + 1])
        rad = np.sqrt(np.float32(-2.0) * np.log(u1)).astype(np.float32)
        out[:, i] = np.sin(v1).astype(np.float32) * rad
        out[:, i + 1] = np.cos(v1).astype(np.float32) * rad
    out = out.reshape(-1)[:n]
    return (out * np.float32(stddev) + np.float32(mean)).astype(np.float32).reshape(shape)


# Known-answer vectors of Philox4x32-10 (Random123 kat_vectors): (counter, key) -> output
KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF),
     (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
     (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
]
