"""numpy restatement of the Philox4x32-10 + Box-Muller sampler used by kernels (1)/(2).

TEST INFRASTRUCTURE.  Philox4x32-10 is the published counter-based generator of Salmon, Moraes,
Dror & Shaw, "Parallel random numbers: as easy as 1, 2, 3" (SC'11) -- the same family
torch.randn_like uses on CUDA -- with the counter/key layout documented in
fd-scorematching_b200/csrc/fdsm_common.cuh.  Known-answer vectors from the Random123 distribution
(kat_vectors: philox4x32 10 rounds) pin the round function in tests/test_oracle_cpu.py.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays.  Returns 4 uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(a, dtype=np.uint32).copy() for a in (c0, c1, c2, c3))
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & MASK).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & MASK).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32((int(k0) + int(W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def normals(B, D, stream_id, seed, offset, row0=0):
    """(B, D) float32 standard normals exactly as fdsm_philox_normal lays them out."""
    nchunk = (D + 3) // 4
    rows = (np.arange(B, dtype=np.uint64) + np.uint64(row0))[:, None].repeat(nchunk, 1)
    chunk = np.arange(nchunk, dtype=np.uint32)[None, :].repeat(B, 0)
    c1 = (rows & MASK).astype(np.uint32)
    c2 = ((rows >> np.uint64(32)).astype(np.uint32) & np.uint32(0x0FFFFFFF)) | np.uint32(stream_id << 28)
    off = ((offset & 0xFFFFFFFF) ^ (((offset >> 32) * 0x9E3779B9) & 0xFFFFFFFF)) & 0xFFFFFFFF
    c3 = np.full_like(c1, np.uint32(off))
    r = philox4x32_10(chunk, c1, c2, c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    k = np.float32(2.0 ** -24)
    u = [((x >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * k for x in r]
    out = np.empty((B, nchunk, 4), dtype=np.float32)
    for p, (ua, ub) in enumerate(((u[0], u[1]), (u[2], u[3]))):
        rad = np.sqrt(np.float32(-2.0) * np.log(ua.astype(np.float64))).astype(np.float32)
        ang = 2.0 * np.pi * ub.astype(np.float64)
        out[:, :, 2 * p] = rad * np.cos(ang).astype(np.float32)
        out[:, :, 2 * p + 1] = rad * np.sin(ang).astype(np.float32)
    return out.reshape(B, nchunk * 4)[:, :D]
