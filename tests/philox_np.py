"""numpy restatement of the engine's random streams (cgnn_b200/csrc/common.cuh): Philox4x32-10, the
(0,1) mapping and the Box-Muller pairing.  The integer part is bit-exact; the float part uses numpy's
float32 log / sqrt / sin / cos, which differ from CUDA's by an ulp or so."""
import numpy as np

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
STREAM_NOISE, STREAM_INIT = 0, 1


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c = [np.asarray(v, np.uint64) & 0xFFFFFFFF for v in (c0, c1, c2, c3)]
    k0, k1 = np.uint64(k0 & 0xFFFFFFFF), np.uint64(k1 & 0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(M0) * c[0]
        p1 = np.uint64(M1) * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & np.uint64(0xFFFFFFFF)
        hi1, lo1 = p1 >> np.uint64(32), p1 & np.uint64(0xFFFFFFFF)
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + np.uint64(W0)) & np.uint64(0xFFFFFFFF)
        k1 = (k1 + np.uint64(W1)) & np.uint64(0xFFFFFFFF)
    return [v.astype(np.uint32) for v in c]


def u01(r):
    return ((r >> np.uint32(9)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 8388608.0)


def normal4(r):
    r0 = np.sqrt(np.float32(-2.0) * np.log(u01(r[0])))
    r1 = np.sqrt(np.float32(-2.0) * np.log(u01(r[2])))
    a0 = np.float32(2.0) * u01(r[1])
    a1 = np.float32(2.0) * u01(r[3])
    pi = np.float64(np.pi)
    return [r0 * np.cos(pi * a0).astype(np.float32), r0 * np.sin(pi * a0).astype(np.float32),
            r1 * np.cos(pi * a1).astype(np.float32), r1 * np.sin(pi * a1).astype(np.float32)]


def pairwise_noise(seed, uid, step, n_rows):
    """e of GNN.py:64 for one sess.run as the pairwise kernel draws it: quad q -> rows 4q..4q+3."""
    q = np.arange((n_rows + 3) // 4)
    n = normal4(philox4x32_10(q, step, STREAM_NOISE, uid, seed & 0xFFFFFFFF, seed >> 32))
    return np.stack(n, 1).reshape(-1)[:n_rows]


def init_params(seed, uid, n_params, init_std):
    q = np.arange((n_params + 3) // 4)
    n = normal4(philox4x32_10(q, 0, STREAM_INIT, uid, seed & 0xFFFFFFFF, seed >> 32))
    return (np.float32(init_std) * np.stack(n, 1).reshape(-1)[:n_params]).astype(np.float32)
