"""CPU: the throughput-mode generator.  (1) oracle/philox.py reproduces the published Philox4x32-10 known-answer vectors
(Random123 kat_vectors).  (2) slam_b200/csrc/slam_rng.cuh — the header the CUDA kernels include — compiled for the host
gives the same words, the same rand()-like ints / resample uniforms bit for bit, and the same normals (to an ulp: the
kernel header multiplies by 1/RAND_MAX where the reference divides) as the numpy twin run through the oracle's so_randn
(core.cpp:383-419).  The device side of the same comparison is tests/test_gpu_round2.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import philox

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = [  # counter, key -> output (Random123 kat_vectors, philox4x32 10 rounds)
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


@pytest.fixture(scope="module")
def emul():
    d = os.path.join(HERE, "host_emul")
    so = os.path.join(d, "librngemul.so")
    src = os.path.join(d, "rng_emul.cpp")
    hdr = os.path.join(HERE, "..", "slam_b200", "csrc", "slam_rng.cuh")
    if (not os.path.exists(so)) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-msse2", "-mfpmath=sse", "-ffp-contract=off", "-fPIC",
                               "-shared", "-x", "c++", "-o", so, src, "-lm"])
    return C.CDLL(so)


def test_philox_known_answers():
    for ctr, key, want in KAT:
        got = philox.philox4x32_10(*[np.array([c], np.uint32) for c in ctr], key[0], key[1])
        assert tuple(int(g[0]) for g in got) == want


def test_kernel_header_matches_known_answers_and_twin(emul, oracle):
    for ctr, key, want in KAT:
        c, k, out = (C.c_uint * 4)(*ctr), (C.c_uint * 2)(*key), (C.c_uint * 4)()
        emul.re_philox(c, k, out)
        assert tuple(out) == want
    seed, n = 0x1234ABCD5678EF01, 4000
    slots = np.arange(n)
    for draw in (0, 1, 77):
        for stream in (0, 1, 2):
            tw = philox.words(seed, slots[:16], draw, stream)
            for p in range(16):
                out = (C.c_uint * 4)()
                emul.re_words(C.c_ulonglong(seed), C.c_ulonglong(p), draw, stream, out)
                assert tuple(out) == tuple(int(w[p]) for w in tw)
        u = np.zeros(n, np.float64)
        emul.re_unif(C.c_ulonglong(seed), C.c_long(n), draw, u.ctypes.data_as(C.c_void_p))
        assert np.array_equal(u, philox.resample_unif(seed, slots, draw))          # bit-exact: keep[] depends on it
        assert u.min() > 0.0 and u.max() <= 1.0
        g3 = np.zeros((3, n), np.float32)
        emul.re_gauss3(C.c_ulonglong(seed), C.c_long(n), draw, g3.ctypes.data_as(C.c_void_p))
        tw3 = philox.pose_gauss(oracle, seed, slots, draw)
        assert np.abs(g3 - tw3).max() <= 2e-6 and np.isfinite(g3).all()
        g2 = np.zeros((2, n), np.float32)
        emul.re_gauss2(C.c_ulonglong(seed), C.c_long(n), draw, g2.ctypes.data_as(C.c_void_p))
        assert np.abs(g2 - philox.control_gauss(oracle, seed, slots, draw)).max() <= 2e-6
    # the normals look like normals
    g = np.zeros((3, 200000), np.float32)
    emul.re_gauss3(C.c_ulonglong(99), C.c_long(200000), 5, g.ctypes.data_as(C.c_void_p))
    assert abs(g.mean()) < 0.01 and abs(g.std() - 1.0) < 0.01 and abs(np.corrcoef(g)[0, 1]) < 0.01
