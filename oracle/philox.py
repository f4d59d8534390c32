"""TEST INFRASTRUCTURE (oracle twin of slam_b200/csrc/slam_rng.cuh) — never imported by the product.

numpy restatement of the throughput-mode generator: Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random
numbers: as easy as 1, 2, 3", SC'11; the constants and round structure below are the published ones, pinned by the
Random123 known-answer vectors in tests/test_rng.py), counter = (global particle lo, hi, draw, stream), key = seed.
A 32-bit word becomes a rand()-like int in [1, RAND_MAX]; from there on the reference's own arithmetic applies:
unifRand() = rand() / double(RAND_MAX) (core.cpp:775-777) and nRandMat::randn's Box-Muller pairing
(core.cpp:383-419, restated in oracle/slam_oracle.c so_randn)."""
import numpy as np

RAND_MAX = 2147483647
STREAM_POSE, STREAM_RESAMPLE, STREAM_CONTROL = 0, 1, 2
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays c0..c3; k0, k1 python ints.  Returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, np.uint64) & _MASK for c in (c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(np.asarray(c, np.uint32) for c in (c0, c1, c2, c3))


def words(seed, slots, draw, stream):
    slots = np.asarray(slots, np.uint64)
    n = slots.shape
    return philox4x32_10(slots & _MASK, slots >> np.uint64(32), np.full(n, draw, np.uint64),
                         np.full(n, stream, np.uint64), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)


def rand_ints(w):
    """The rand()-like draw of one word: word >> 1, with 0 mapped to 1."""
    return np.maximum((np.asarray(w, np.uint32) >> np.uint32(1)).astype(np.int32), 1)


def resample_unif(seed, slots, draw):
    """[n] doubles: the uniforms stratifiedRandom consumes on update number `draw`."""
    return rand_ints(words(seed, slots, draw, STREAM_RESAMPLE)[0]) / float(RAND_MAX)


def resample_ints(seed, slots, draw):
    return rand_ints(words(seed, slots, draw, STREAM_RESAMPLE)[0])


def _randn(oracle, ints, count):
    """ints [count + 1][n] -> [count][n] standard normals through the oracle's so_randn."""
    n = ints.shape[1]
    out = np.empty((count, n), np.float32)
    cols = np.ascontiguousarray(ints.T, np.int32)
    for p in range(n):
        out[:, p] = oracle.randn(cols[p], count)
    return out


def pose_gauss(oracle, seed, slots, draw):
    """[3][n] normals of the FastSLAM2 pose sample of update number `draw`."""
    w = words(seed, slots, draw, STREAM_POSE)
    return _randn(oracle, np.stack([rand_ints(x) for x in w]), 3)


def control_gauss(oracle, seed, slots, draw):
    """[2][n] normals of the control noise of predict tick number `draw`."""
    w = words(seed, slots, draw, STREAM_CONTROL)
    return _randn(oracle, np.stack([rand_ints(x) for x in w[:3]]), 2)
