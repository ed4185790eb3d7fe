"""Seeded ray batches for the traversal configuration (C4): the same rays here (oracle) and on
the GPU box.  Rays come in blocks of 2^16; block b is drawn from Philox(key=SEED, counter=b), so
any sub-range can be produced without generating what precedes it.

Incoherent rays: origins uniform in the scene bound, directions uniform on the sphere
(UniformSampleSphere form), tMax = inf -- the scheme of src/tests/shapes.cpp:98-109.
"""
import numpy as np

SEED = 20240611
BLOCK = 1 << 16
RAY = np.dtype([("o", "<f4", 3), ("t_max", "<f4"), ("d", "<f4", 3), ("time", "<f4")])


def _block(b, lo, hi):
    g = np.random.Generator(np.random.Philox(key=SEED, counter=b))
    u = g.random((BLOCK, 5), dtype=np.float32)
    lo = np.asarray(lo, np.float32)
    ext = np.asarray(hi, np.float32) - lo
    rays = np.zeros(BLOCK, dtype=RAY)
    rays["o"] = lo + u[:, 0:3] * ext
    z = 1 - 2 * u[:, 3]
    r = np.sqrt(np.maximum(0, 1 - z * z))
    phi = np.float32(2 * np.pi) * u[:, 4]
    rays["d"] = np.stack([r * np.cos(phi), r * np.sin(phi), z], -1).astype(np.float32)
    rays["t_max"] = np.inf
    return rays


def incoherent_rays(n, lo, hi, first=0):
    assert first % BLOCK == 0
    out = np.zeros(n, dtype=RAY)
    b, k = first // BLOCK, 0
    while k < n:
        blk = _block(b, lo, hi)
        m = min(BLOCK, n - k)
        out[k:k + m] = blk[:m]
        k += m
        b += 1
    return out
