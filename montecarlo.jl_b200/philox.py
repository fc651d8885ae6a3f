"""Vectorised Philox4x32-10 (numpy) for host-side initial configurations.

The engine's acceptance uniforms are generated on the device (csrc/philox.cuh);
this module draws the *inputs* the reference would draw with `rand(...)`
(src/models/HubbardModel/HubbardModel.jl:46-48, src/models/Ising/IsingModel.jl:72)
from the same counter-based family, keyed by the GLOBAL chain id so results do
not depend on how chains are sharded over GPUs.
"""
import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr, key):
    """ctr: (..., 4) uint32, key: (..., 2) uint32 -> (..., 4) uint32."""
    ctr = np.asarray(ctr, dtype=np.uint64)
    key = np.asarray(key, dtype=np.uint64)
    c0, c1, c2, c3 = (ctr[..., i].copy() for i in range(4))
    k0, k1 = key[..., 0].copy(), key[..., 1].copy()
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        n0 = (p1 >> np.uint64(32)) ^ c1 ^ k0
        n1 = p1 & _MASK
        n2 = (p0 >> np.uint64(32)) ^ c3 ^ k1
        n3 = p0 & _MASK
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(_W0)) & _MASK
        k1 = (k1 + np.uint64(_W1)) & _MASK
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def seed_key(seed, stream):
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    return np.array([(seed & 0xFFFFFFFF) ^ (seed >> 32), int(stream) & 0xFFFFFFFF], dtype=np.uint32)


def random_bits(n, seed, stream, domain):
    """n pseudo-random bits for (seed, stream); `domain` separates uses (counter word 0)."""
    groups = (n + 127) // 128
    ctr = np.zeros((groups, 4), dtype=np.uint32)
    ctr[:, 0] = domain
    g = np.arange(groups, dtype=np.uint64)
    ctr[:, 2] = (g & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ctr[:, 3] = (g >> np.uint64(32)).astype(np.uint32)
    words = philox4x32_10(ctr, np.broadcast_to(seed_key(seed, stream), (groups, 2)))
    bits = np.unpackbits(words.view(np.uint8).reshape(groups, 16), axis=1, bitorder="little")
    return bits.reshape(-1)[:n]


def dqmc_uniform(seed, chain, sweep, visit, site):
    """Acceptance uniform of the DQMC sweep (mirror of csrc/philox.cuh dqmc_uniform)."""
    site = np.atleast_1d(np.asarray(site, dtype=np.uint32))
    ctr = np.zeros((site.size, 4), dtype=np.uint32)
    ctr[:, 0] = sweep
    ctr[:, 1] = visit
    ctr[:, 2] = site
    w = philox4x32_10(ctr, np.broadcast_to(seed_key(seed, chain), (site.size, 2))).astype(np.uint64)
    x = (w[:, 1] << np.uint64(32)) | w[:, 0]
    return (x >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
