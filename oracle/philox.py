"""Numpy restatement of the production draw contract (TEST INFRASTRUCTURE: only tests/,
smoke() and bench.py's CPU legs may import this package).

The reference draws from numpy's global Mersenne Twister (``np.random.normal(size=d)``,
``np.random.uniform`` -- src/pyhmc.py:66,59); the GPU path cannot, and uses the counter-based
Philox4x32-10 of Salmon et al. (SC'11, the Random123 library) instead.  This file restates
that generator from its published definition and the way libhmcb keys it, so that a test can
compute on the host exactly which uniforms -- and, to the accuracy of the hardware's
log2/sin/cos approximations, which normals -- a chain consumed in a production run, and replay
them through the reference's transition (``oracle/samplers.py``, pinned bit-exact to the
unmodified reference classes).

  key     = (seed lo, seed hi)
  counter = (chain lo, chain hi, iteration, stream << 24 | block)
  streams : 0 momentum normals, 1 accept-test uniforms, 2 step-size draws
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
STREAM_MOMENTUM, STREAM_ACCEPT, STREAM_STEPSIZE = 0, 1, 2


def philox4x32_10(ctr, key):
    """ctr: uint32[..., 4]; key: (k0, k1) scalars or uint32[...] arrays.  Ten rounds of
    (hi, lo) = M * c multiplies with the Weyl key schedule."""
    c = np.asarray(ctr, dtype=np.uint64)
    c0, c1, c2, c3 = c[..., 0], c[..., 1], c[..., 2], c[..., 3]
    k0 = np.asarray(key[0], dtype=np.uint64)
    k1 = np.asarray(key[1], dtype=np.uint64)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2          # < 2^64: no overflow in uint64
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0 = (k0 + np.uint64(W0)) & MASK
        k1 = (k1 + np.uint64(W1)) & MASK
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def block(seed, chain, iteration, stream, blk):
    ctr = np.array([chain & 0xFFFFFFFF, chain >> 32, iteration & 0xFFFFFFFF, ((stream << 24) | (blk & 0xFFFFFF))],
                   dtype=np.uint64)
    return philox4x32_10(ctr, (seed & 0xFFFFFFFF, seed >> 32))


def blocks(seed, chains, iteration, stream, n_blocks, first_block=0):
    """uint32[C, n_blocks, 4] for global chain ids ``chains``."""
    chains = np.asarray(chains, dtype=np.uint64)
    C = chains.shape[0]
    ctr = np.empty((C, n_blocks, 4), dtype=np.uint64)
    ctr[..., 0] = (chains & MASK)[:, None]
    ctr[..., 1] = (chains >> np.uint64(32))[:, None]
    ctr[..., 2] = np.uint64(iteration & 0xFFFFFFFF)
    ctr[..., 3] = (np.uint64(stream << 24) | (np.arange(first_block, first_block + n_blocks, dtype=np.uint64)
                                               & np.uint64(0xFFFFFF)))[None, :]
    return philox4x32_10(ctr, (seed & 0xFFFFFFFF, seed >> 32))


def u53(a, b):
    """53-bit uniform in [0, 1): the top 27 bits of ``a`` over the top 26 bits of ``b``."""
    a = np.asarray(a, dtype=np.uint64)
    b = np.asarray(b, dtype=np.uint64)
    return (((a >> np.uint64(5)) << np.uint64(26)) | (b >> np.uint64(6))).astype(np.float64) * (1.0 / 9007199254740992.0)


def accept_uniforms(seed, chains, iteration, n):
    """float64[C, n]: uniform i of a chain = words (0,1) of accept-stream block i//2 for even i, words (2,3) for odd."""
    w = blocks(seed, chains, iteration, STREAM_ACCEPT, (n + 1) // 2)
    u = np.empty((w.shape[0], 2 * w.shape[1]))
    u[:, 0::2] = u53(w[..., 0], w[..., 1])
    u[:, 1::2] = u53(w[..., 2], w[..., 3])
    return u[:, :n]


def momentum_normals_f32(seed, chains, iteration, d):
    """float64[C, d]: the f32 Box-Muller of the kernels evaluated with correctly rounded float32
    log2 / sin / cos.  The kernels use the hardware approximations (MUFU), so a drawn momentum
    agrees with this to ~1e-6 absolute, not bit for bit."""
    nb = (d + 3) // 4
    w = blocks(seed, chains, iteration, STREAM_MOMENTUM, nb).astype(np.float32)          # [C, nb, 4]
    z = np.empty((w.shape[0], nb, 4), dtype=np.float32)
    for h in range(2):
        u1 = (w[..., 2 * h] + np.float32(0.5)) * np.float32(2.3283064365386963e-10)
        r = np.sqrt(np.float32(-1.3862943611198906) * np.log2(u1, dtype=np.float32), dtype=np.float32)
        th = w[..., 2 * h + 1] * np.float32(1.4629180792671596e-09)
        z[..., 2 * h] = r * np.cos(th, dtype=np.float32)
        z[..., 2 * h + 1] = r * np.sin(th, dtype=np.float32)
    return z.reshape(w.shape[0], nb * 4)[:, :d].astype(np.float64)


def momentum_normals_f64(seed, chains, iteration, d):
    """float64[C, d]: the HMCB_RNG_NORMAL_F64 generator -- element block b takes Philox blocks 2b and 2b+1, each
    giving two normals from 53-bit uniforms."""
    nb = (d + 3) // 4
    w = blocks(seed, chains, iteration, STREAM_MOMENTUM, 2 * nb)                          # [C, 2nb, 4]
    u1 = 1.0 - u53(w[..., 0], w[..., 1])
    r = np.sqrt(-2.0 * np.log(u1))
    ang = 2.0 * u53(w[..., 2], w[..., 3])
    z = np.stack([r * np.cos(np.pi * ang), r * np.sin(np.pi * ang)], axis=-1)             # [C, 2nb, 2]
    return z.reshape(w.shape[0], 4 * nb)[:, :d]
