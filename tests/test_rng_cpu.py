"""The draw contract of the production path, pinned on the host (no GPU): the library's
Philox4x32-10 against the published Random123 known-answer vectors (Salmon, Moraes, Dror,
Shaw, "Parallel random numbers: as easy as 1, 2, 3", SC'11; kat_vectors of Random123), the
counter layout the kernels use, and the 53-bit uniform against a numpy restatement.  The
kernels call the very same ``Philox::gen`` / ``u53`` (``__host__ __device__`` in
csrc/common.cuh); tests/test_gpu_bench_path.py checks on the GPU that the momenta and
uniforms a chain consumes are these."""
import ctypes as C

import numpy as np

from hmc_b200 import _lib
from oracle import philox as ophilox

U32x4 = C.c_uint32 * 4
U32x2 = C.c_uint32 * 2

# Random123 kat_vectors: philox4x32-10   counter[4]   key[2]   expected[4]
KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff),
     (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def _lib_block(ctr, key):
    out = U32x4()
    _lib.load().hmcb_philox4x32_10(U32x4(*ctr), U32x2(*key), out)
    return tuple(out)


def test_philox4x32_10_known_answers():
    for ctr, key, want in KAT:
        assert _lib_block(ctr, key) == want, "library Philox differs from Random123 at counter %r" % (ctr,)
        assert tuple(int(x) for x in ophilox.philox4x32_10(np.array(ctr, dtype=np.uint32), key)) == want, \
            "oracle Philox differs from Random123"


def test_counter_layout_of_a_kernel_draw():
    """key = seed (lo, hi); counter = (chain lo, chain hi, iteration, stream << 24 | block)."""
    lib = _lib.load()
    rng = np.random.default_rng(0)
    for _ in range(50):
        seed = int(rng.integers(0, 2 ** 63)) * 2 + int(rng.integers(0, 2))
        chain = int(rng.integers(0, 2 ** 40))
        it, stream, block = int(rng.integers(0, 2 ** 32)), int(rng.integers(0, 3)), int(rng.integers(0, 2 ** 24))
        out = U32x4()
        lib.hmcb_philox_block(seed, chain, it, stream, block, out)
        ctr = (chain & 0xffffffff, chain >> 32, it, (stream << 24) | block)
        key = (seed & 0xffffffff, seed >> 32)
        assert tuple(out) == _lib_block(ctr, key)
        assert tuple(out) == tuple(int(x) for x in ophilox.block(seed, chain, it, stream, block))


def test_u53_matches_numpy_restatement():
    lib = _lib.load()
    rng = np.random.default_rng(1)
    w = rng.integers(0, 2 ** 32, size=(1000, 2), dtype=np.uint64)
    w[0], w[1], w[2] = (0, 0), (2 ** 32 - 1, 2 ** 32 - 1), (1 << 5, 1 << 6)
    for a, b in w:
        got = lib.hmcb_u53(int(a), int(b))
        want = float(ophilox.u53(np.uint32(a), np.uint32(b)))
        assert got == want and 0.0 <= got < 1.0
    assert lib.hmcb_u53(2 ** 32 - 1, 2 ** 32 - 1) == 1.0 - 2.0 ** -53


def test_vectorised_oracle_philox_equals_library():
    """oracle/philox.py generates whole batches (chains x blocks) at once for the GPU tests; it must be the
    library's generator word for word."""
    lib = _lib.load()
    chains = np.arange(1000, 1017, dtype=np.uint64)
    got = ophilox.blocks(seed=123456789012345, chains=chains, iteration=7, stream=0, n_blocks=13)   # [C, 13, 4]
    for ci, ch in enumerate(chains):
        for b in range(13):
            out = U32x4()
            lib.hmcb_philox_block(123456789012345, int(ch), 7, 0, b, out)
            assert tuple(out) == tuple(int(x) for x in got[ci, b])
    u = ophilox.accept_uniforms(seed=5, chains=chains, iteration=3, n=5)                               # [C, 5]
    for ci, ch in enumerate(chains):
        for i in range(5):
            out = U32x4()
            lib.hmcb_philox_block(5, int(ch), 3, 1, i >> 1, out)
            want = lib.hmcb_u53(out[2], out[3]) if i & 1 else lib.hmcb_u53(out[0], out[1])
            assert u[ci, i] == want
