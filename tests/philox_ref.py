"""Independent numpy Philox4x32-10 (Salmon et al. SC'11) used by the tests to predict
the device's random choices.  Checked against the Random123 known-answer vectors in
tests/test_host_logic.py."""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c = [np.asarray(x, dtype=np.uint64) & MASK for x in (c0, c1, c2, c3)]
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c = [(hi1 ^ c[1] ^ np.uint64(k0)) & MASK, lo1, (hi0 ^ c[3] ^ np.uint64(k1)) & MASK, lo0]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return c


def draw(seed, gid, stream, ctr):
    """The engine's philox_draw(seed, gid, stream, ctr) -> uint32 (vectorised over gid / ctr)."""
    gid = np.asarray(gid, dtype=np.uint64)
    ctr = np.asarray(ctr, dtype=np.uint64)
    blk = philox4x32_10(ctr >> np.uint64(2), np.full_like(gid, stream), gid & MASK, gid >> np.uint64(32),
                        seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    sel = (ctr & np.uint64(3)).astype(np.int64)
    out = np.choose(sel, [np.broadcast_to(b, sel.shape) for b in blk])
    return out.astype(np.uint64)


def bounded(r, n):
    """Lemire multiply-shift used on the device: floor(r * n / 2^32)"""
    return ((np.asarray(r, dtype=np.uint64) * np.asarray(n, dtype=np.uint64)) >> np.uint64(32)).astype(np.int64)
