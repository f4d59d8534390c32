"""TEST INFRASTRUCTURE ONLY — ctypes loaders for the two CPU checkers.

* ``Oracle``  : oracle/liboracle.so, the plain-C restatement (oracle/slam_oracle.c).
* ``Ref``     : oracle/_ref/libslamref.so, the UNMODIFIED reference sources compiled in place from
                /root/reference plus oracle/ref_shim.cpp (prebuilt; never rebuilt on the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package (slam_b200) never does.

Flat layout shared with include/slamgpu.h:
  xv[3][P] | Pv[9][P] (row-major 3x3 planes) | w[P] | lm_mean[Nf][P][2] | lm_cov[Nf][P][4] (column-major 2x2)
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
RAND_MAX = 2147483647

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def build(ref=True):
    """(Re)build liboracle.so and, when /root/reference exists, _ref/libslamref.so."""
    subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"] + (["ref"] if ref else []))


class State:
    """Flat particle set (numpy) in the shared layout."""

    def __init__(self, P, nf_cap):
        self.P, self.nf_cap, self.Nf = P, nf_cap, 0
        self.xv = np.zeros((3, P), np.float32)
        self.Pv = np.zeros((9, P), np.float32)
        self.w = np.full(P, np.float32(1.0 / np.float32(P)), np.float32)
        self.lm_mean = np.zeros((nf_cap, P, 2), np.float32)
        self.lm_cov = np.zeros((nf_cap, P, 4), np.float32)

    def copy(self):
        s = State.__new__(State)
        s.P, s.nf_cap, s.Nf = self.P, self.nf_cap, self.Nf
        for k in ("xv", "Pv", "w", "lm_mean", "lm_cov"):
            setattr(s, k, getattr(self, k).copy())
        return s


class Oracle:
    def __init__(self):
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = self.lib = C.CDLL(path)
        L.so_trig_offset.restype = C.c_float
        L.so_trig_offset.argtypes = [C.c_float]
        L.so_randn.restype = C.c_int
        L.so_randn.argtypes = [i32p, C.c_int, f32p]
        L.so_eigen_sum.restype = C.c_float
        L.so_eigen_sum.argtypes = [f32p, C.c_long]
        L.so_compute_jacobians.argtypes = [f32p, C.c_int, f32p]
        L.so_gauss_evaluate.restype = C.c_float
        L.so_gauss_evaluate.argtypes = [f32p, f32p, C.c_int]
        L.so_cholesky_update.argtypes = [f32p, f32p, f32p, f32p, f32p]
        L.so_joseph_update.argtypes = [f32p, f32p, C.c_float, C.c_float]
        L.so_multivariate_gauss.argtypes = [C.c_int, f32p, f32p, f32p, f32p]
        pred = [C.c_long, f32p, f32p, C.c_float, C.c_float, f32p, C.c_float, C.c_float, C.c_float, C.c_float,
                C.c_int, C.c_void_p]
        L.so_fs2_predict.argtypes = pred
        L.so_fs1_predict.argtypes = pred
        L.so_fs2_update_particles.restype = C.c_int
        L.so_fs2_update_particles.argtypes = [C.c_long, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, C.c_int,
                                              f32p, C.c_int, f32p, f32p]
        L.so_fs1_update_particles.restype = C.c_int
        L.so_fs1_update_particles.argtypes = [C.c_long, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, C.c_int,
                                              f32p, C.c_int, f32p]
        L.so_stratified_count.restype = C.c_long
        L.so_stratified_count.argtypes = [C.c_long]
        L.so_stratified_di.argtypes = [C.c_long, f32p]
        L.so_resample_indices.argtypes = [C.c_long, f32p, f64p, C.c_int, i32p, f32p, C.c_void_p]
        L.so_resample_gather.argtypes = [C.c_long, C.c_int, i32p, f32p, f32p, f32p, f32p, f32p]
        L.so_estimate.argtypes = [C.c_long, f32p, f32p, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                  C.POINTER(C.c_double)]

    # -- primitives -------------------------------------------------------
    def trig_offset(self, a):
        return self.lib.so_trig_offset(float(a))

    def randn(self, rnd, count):
        out = np.zeros(count, np.float32)
        self.lib.so_randn(_i32(rnd), count, out)
        return out

    def eigen_sum(self, v):
        v = _f32(v)
        return self.lib.so_eigen_sum(v, v.size)

    def compute_jacobians(self, rec, n):
        out = np.zeros(16 * n, np.float32)
        self.lib.so_compute_jacobians(_f32(rec), n, out)
        return out

    def gauss_evaluate(self, v, S):
        v = _f32(v)
        return self.lib.so_gauss_evaluate(v, _f32(S).ravel(), v.size)

    def cholesky_update(self, x, Plin, v, R, H):
        x, Plin = _f32(x).copy(), _f32(Plin).copy()
        self.lib.so_cholesky_update(x, Plin, _f32(v), _f32(R).ravel(), _f32(H).ravel())
        return x, Plin

    def joseph_update(self, x, P, v, R):
        x, P = _f32(x).copy(), _f32(P).ravel().copy()
        self.lib.so_joseph_update(x, P, float(v), float(R))
        return x, P.reshape(3, 3)

    def multivariate_gauss(self, x, P, g):
        x = _f32(x)
        out = np.zeros(x.size, np.float32)
        self.lib.so_multivariate_gauss(x.size, x, _f32(P).ravel(), _f32(g), out)
        return out

    # -- particle-set steps (in place on a State) --------------------------
    def predict(self, method, s, V, G, Q, dt, phi_true, wheel_base, sigma_phi, use_heading, gauss=None):
        fn = self.lib.so_fs2_predict if method == 2 else self.lib.so_fs1_predict
        g = None
        if gauss is not None:
            g = _f32(gauss)
            assert g.shape == (2, s.P)
        fn(s.P, s.xv, s.Pv, V, G, _f32(Q).ravel(), dt, phi_true, wheel_base, sigma_phi, int(use_heading),
           None if g is None else g.ctypes.data_as(C.c_void_p))

    def update_particles(self, method, s, zf, idf, zn, R, gauss=None):
        zf = _f32(zf).reshape(-1, 2)
        zn = _f32(zn).reshape(-1, 2)
        idf = _i32(idf)
        K, Kn = zf.shape[0], zn.shape[0]
        assert s.Nf + Kn <= s.nf_cap
        zf_ = zf if K else np.zeros((1, 2), np.float32)
        zn_ = zn if Kn else np.zeros((1, 2), np.float32)
        idf_ = idf if K else np.zeros(1, np.int32)
        if method == 2:
            g = _f32(gauss)
            assert g.shape == (3, s.P)
            s.Nf = self.lib.so_fs2_update_particles(s.P, s.Nf, s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov, zf_, idf_, K,
                                                    zn_, Kn, _f32(R).ravel(), g)
        else:
            s.Nf = self.lib.so_fs1_update_particles(s.P, s.Nf, s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov, zf_, idf_, K,
                                                    zn_, Kn, _f32(R).ravel())

    def stratified_count(self, N):
        return self.lib.so_stratified_count(N)

    def stratified_di(self, N):
        di = np.zeros(N, np.float32)
        self.lib.so_stratified_di(N, di)
        return di

    def resample_indices(self, w, unif, scan_mode=0, want_cdf=False):
        """Returns (normalised w, keep, nEff[, cdf])."""
        w = _f32(w).copy()
        P = w.size
        unif = np.ascontiguousarray(unif, np.float64)
        keep = np.zeros(P, np.int32)
        neff = np.zeros(1, np.float32)
        cdf = np.zeros(P, np.float32) if want_cdf else None
        self.lib.so_resample_indices(P, w, unif, scan_mode, keep, neff,
                                     None if cdf is None else cdf.ctypes.data_as(C.c_void_p))
        return (w, keep, neff[0], cdf) if want_cdf else (w, keep, neff[0])

    def resample_gather(self, s, keep):
        self.lib.so_resample_gather(s.P, s.Nf, _i32(keep), s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov)

    def resample(self, s, unif, n_effective, do_resample=True, scan_mode=0):
        """resampleParticles (core.cpp:718-749) on a State.  Returns (keep, nEff, did_resample)."""
        wn, keep, neff = self.resample_indices(s.w, unif, scan_mode)
        s.w[:] = wn
        did = bool(do_resample and (neff < n_effective))
        if did:
            self.resample_gather(s, keep)
        return keep, neff, did

    def estimate(self, s):
        x, y, t = C.c_double(), C.c_double(), C.c_double()
        self.lib.so_estimate(s.P, s.xv, s.w, C.byref(x), C.byref(y), C.byref(t))
        return x.value, y.value, t.value


class Ref:
    """The unmodified reference, through oracle/ref_shim.cpp."""

    @staticmethod
    def path():
        return os.path.join(HERE, "_ref", "libslamref.so")

    @staticmethod
    def available():
        return os.path.exists(Ref.path())

    def __init__(self):
        L = self.lib = C.CDLL(Ref.path())
        L.slamref_rand_set_stream.argtypes = [i32p, C.c_long]
        L.slamref_rand_pos.restype = C.c_long
        L.slamref_rand_underflow.restype = C.c_long
        L.slamref_randn.argtypes = [C.c_int, C.c_int, f32p]
        L.slamref_trig_offset.restype = C.c_float
        L.slamref_trig_offset.argtypes = [C.c_float]
        L.slamref_multivariate_gauss.argtypes = [C.c_int, f32p, f32p, f32p]
        L.slamref_sum.restype = C.c_float
        L.slamref_sum.argtypes = [f32p, C.c_int]
        L.slamref_cumsum.argtypes = [f32p, C.c_int]
        L.slamref_stratified_count.restype = C.c_long
        L.slamref_stratified_count.argtypes = [C.c_long]
        L.slamref_stratified_random.argtypes = [C.c_long, f32p]
        L.slamref_stratified_resample.argtypes = [f32p, C.c_int, i32p, f32p]
        L.slamref_compute_jacobians.argtypes = [f32p, C.c_int, f32p]
        L.slamref_gauss_evaluate.restype = C.c_float
        L.slamref_gauss_evaluate.argtypes = [f32p, f32p, C.c_int]
        L.slamref_cholesky_update.argtypes = [f32p, f32p, f32p, f32p, f32p]
        L.slamref_joseph_update.argtypes = [f32p, f32p, C.c_float, C.c_float]
        L.slamref_create.restype = C.c_void_p
        L.slamref_create.argtypes = [C.c_int]
        L.slamref_destroy.argtypes = [C.c_void_p]
        L.slamref_num_landmarks.argtypes = [C.c_void_p]
        L.slamref_upload.argtypes = [C.c_void_p, C.c_int, f32p, f32p, f32p, f32p, f32p]
        L.slamref_download.argtypes = [C.c_void_p, f32p, f32p, f32p, f32p, f32p]
        pred = [C.c_void_p, C.c_float, C.c_float, f32p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int]
        L.slamref_fs2_predict.argtypes = pred
        L.slamref_fs1_predict.argtypes = pred
        upd = [C.c_void_p, f32p, i32p, C.c_int, f32p, C.c_int, f32p]
        L.slamref_fs2_update.argtypes = upd + [C.c_int, C.c_int]
        L.slamref_fs1_update.argtypes = upd + [C.c_int, C.c_int]
        L.slamref_fs2_update_particles.argtypes = upd
        L.slamref_fs1_update_particles.argtypes = upd
        L.slamref_resample.argtypes = [C.c_void_p, C.c_int, C.c_int]

    # -- rand() stream ----------------------------------------------------
    def set_stream(self, ints):
        self.lib.slamref_rand_set_stream(_i32(ints), len(ints))

    def underflow(self):
        return self.lib.slamref_rand_underflow()

    def rand_pos(self):
        return self.lib.slamref_rand_pos()

    # -- primitives -------------------------------------------------------
    def randn(self, m, n):
        out = np.zeros(m * n, np.float32)
        self.lib.slamref_randn(m, n, out)
        return out.reshape(m, n)

    def trig_offset(self, a):
        return self.lib.slamref_trig_offset(float(a))

    def multivariate_gauss(self, x, P):
        x = _f32(x)
        out = np.zeros(x.size, np.float32)
        self.lib.slamref_multivariate_gauss(x.size, x, _f32(P).ravel(), out)
        return out

    def sum(self, w):
        w = _f32(w)
        return self.lib.slamref_sum(w, w.size)

    def cumsum(self, w):
        w = _f32(w).copy()
        self.lib.slamref_cumsum(w, w.size)
        return w

    def stratified_count(self, N):
        return self.lib.slamref_stratified_count(N)

    def stratified_random(self, N):
        di = np.zeros(N, np.float32)
        self.lib.slamref_stratified_random(N, di)
        return di

    def stratified_resample(self, w):
        w = _f32(w)
        assert self.stratified_count(w.size) == w.size, "reference asserts at core.cpp:762 for this N"
        keep = np.zeros(w.size, np.int32)
        neff = np.zeros(1, np.float32)
        self.lib.slamref_stratified_resample(w, w.size, keep, neff)
        return keep, neff[0]

    def compute_jacobians(self, rec, n):
        out = np.zeros(16 * n, np.float32)
        self.lib.slamref_compute_jacobians(_f32(rec), n, out)
        return out

    def gauss_evaluate(self, v, S):
        v = _f32(v)
        return self.lib.slamref_gauss_evaluate(v, _f32(S).ravel(), v.size)

    def cholesky_update(self, x, Plin, v, R, H):
        x, Plin = _f32(x).copy(), _f32(Plin).copy()
        self.lib.slamref_cholesky_update(x, Plin, _f32(v), _f32(R).ravel(), _f32(H).ravel())
        return x, Plin

    def joseph_update(self, x, P, v, R):
        x, P = _f32(x).copy(), _f32(P).ravel().copy()
        self.lib.slamref_joseph_update(x, P, float(v), float(R))
        return x, P.reshape(3, 3)

    # -- particle sets ------------------------------------------------------
    def create(self, s):
        h = self.lib.slamref_create(s.P)
        self.upload(h, s)
        return h

    def destroy(self, h):
        self.lib.slamref_destroy(h)

    def upload(self, h, s):
        self.lib.slamref_upload(h, s.Nf, s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov)

    def download(self, h, s):
        s.Nf = self.lib.slamref_num_landmarks(h)
        assert s.Nf <= s.nf_cap
        self.lib.slamref_download(h, s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov)

    def predict(self, method, h, V, G, Q, dt, phi_true, wheel_base, sigma_phi, use_heading, add_noise):
        fn = self.lib.slamref_fs2_predict if method == 2 else self.lib.slamref_fs1_predict
        fn(h, V, G, _f32(Q).ravel(), dt, phi_true, wheel_base, sigma_phi, int(use_heading), int(add_noise))

    @staticmethod
    def _obs(zf, idf, zn):
        zf = _f32(zf).reshape(-1, 2)
        zn = _f32(zn).reshape(-1, 2)
        idf = _i32(idf)
        K, Kn = zf.shape[0], zn.shape[0]
        return (zf if K else np.zeros((1, 2), np.float32), idf if K else np.zeros(1, np.int32), K,
                zn if Kn else np.zeros((1, 2), np.float32), Kn)

    def update_particles(self, method, h, zf, idf, zn, R):
        zf, idf, K, zn, Kn = self._obs(zf, idf, zn)
        fn = self.lib.slamref_fs2_update_particles if method == 2 else self.lib.slamref_fs1_update_particles
        fn(h, zf, idf, K, zn, Kn, _f32(R).ravel())

    def update(self, method, h, zf, idf, zn, R, n_effective, resample):
        """Whole FastSLAM{1,2}::update including resampleParticles."""
        zf, idf, K, zn, Kn = self._obs(zf, idf, zn)
        fn = self.lib.slamref_fs2_update if method == 2 else self.lib.slamref_fs1_update
        fn(h, zf, idf, K, zn, Kn, _f32(R).ravel(), int(n_effective), int(resample))

    def resample(self, h, n_min, do_resample):
        self.lib.slamref_resample(h, int(n_min), int(do_resample))


def randn_from_ints(ints, count):
    """Standard normals the reference's nRandMat::randn(count,1) derives from count+1 raw rand() values."""
    return Oracle().randn(np.asarray(ints[:count + 1], np.int32), count)
