"""Host-side mirror of the reference's FastSLAM method objects over the C ABI (include/slamgpu.h).

The reference drives `FastSLAM2::predict(particles, xTrue, V, G, Q, dt)` and
`FastSLAM2::update(particles, zf, zn, idf, z, dataAssociationTable, R)` (algorithms/fastslam2.h:20-24; FastSLAM1
alike, fastslam1.h:29-33) from its wrapper loop (wrappers/fastslam2wrapper.cpp:64,88).  `FastSLAM` below keeps those
entry points, argument meanings and public switches (addPredictNoise, useHeading, wheelBase, nEffective, resample,
sigmaPhi — fastslam2.h:26-31); `particles` is the device-resident set owned by the handle.  Noise the reference draws
from rand() is passed in explicitly (gauss / unif) so that runs can be replayed against the reference.

This module only marshals arguments; all arithmetic runs in the CUDA library.
"""
import ctypes as C

import numpy as np

from . import _lib

FASTSLAM1, FASTSLAM2 = 1, 2
SCAN_REFERENCE, SCAN_BLOCKED = 0, 1
MATH_FAST, MATH_REFERENCE_ORDER = 0, 1
T_PREDICT, T_UPDATE, T_ADDFEAT, T_WEIGHTS, T_GATHER, T_ESTIMATE = range(6)
T_NAMES = ["predict", "update", "add_feature", "weights", "gather", "estimate"]


class SlamGpuError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("slamgpu error %d: %s" % (code, text))
        self.code = code


def _ptr(a):
    return None if a is None else a.ctypes.data


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class FastSLAM:
    """One FastSLAM1/FastSLAM2 particle set on one CUDA device."""

    def __init__(self, method, n_particles, nf_cap, wheel_base=4.0, sigma_phi=0.0, n_effective=None, use_heading=False,
                 resample=True, scan_mode=SCAN_BLOCKED, device=0, spare_rows=0, math_mode=MATH_FAST, log_epochs=0,
                 keep_nan_weights=False):
        self._L = _lib.load()
        self._h = C.c_void_p()
        self.P = int(n_particles)
        self.nf_cap = int(nf_cap)
        self.method = int(method)
        cfg = _lib.Config()
        cfg.n_particles = self.P
        cfg.nf_cap = self.nf_cap
        cfg.method = self.method
        cfg.wheel_base = wheel_base
        cfg.sigma_phi = sigma_phi
        cfg.n_effective = int(0.75 * self.P) if n_effective is None else int(n_effective)
        cfg.use_heading = int(bool(use_heading))
        cfg.do_resample = int(bool(resample))
        cfg.scan_mode = scan_mode
        cfg.device = device
        cfg.spare_rows = spare_rows
        cfg.math_mode = math_mode
        cfg.log_epochs = log_epochs
        cfg.keep_nan_weights = int(bool(keep_nan_weights))
        self._check(self._L.slamgpu_create(C.byref(cfg), C.byref(self._h)))

    # -- plumbing ---------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise SlamGpuError(rc, self._L.slamgpu_last_error().decode())

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._L.slamgpu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        self._check(self._L.slamgpu_set_stream(self._h, C.c_void_p(cuda_stream)))

    def synchronize(self):
        self._check(self._L.slamgpu_synchronize(self._h))

    @property
    def Nf(self):
        return self._L.slamgpu_landmark_count(self._h)

    @property
    def launches(self):
        return self._L.slamgpu_launch_count(self._h)

    def stats(self):
        """rows re-columned by deferred resampling, resamples applied, ancestor-chain steps."""
        return dict(rows_moved=self._L.slamgpu_stat(self._h, 0), resamples=self._L.slamgpu_stat(self._h, 1),
                    chain_steps=self._L.slamgpu_stat(self._h, 2), rows_fused=self._L.slamgpu_stat(self._h, 3),
                    remote_sources=self._L.slamgpu_stat(self._h, 4), keep_fallbacks=self._L.slamgpu_stat(self._h, 5),
                    weight_resets=self._L.slamgpu_stat(self._h, 6))

    def seed(self, seed, predict_noise=False):
        """Throughput mode: noise the reference draws inside predict/update comes from the device generator."""
        self._check(self._L.slamgpu_seed(self._h, int(seed), int(bool(predict_noise))))

    def debug_capture(self, on=True):
        self._check(self._L.slamgpu_debug_capture(self._h, int(on)))

    def debug_read(self, which):
        """0: weights entering the last resample, 1: after normalisation, 2: keep[], 3: select[]."""
        out = np.empty(self.P, np.int32 if which == 2 else np.float32)
        self._check(self._L.slamgpu_debug_read(self._h, int(which), out.ctypes.data))
        return out

    def download_poses_at(self, idx):
        idx = np.ascontiguousarray(idx, np.int64)
        n = idx.size
        out = dict(xv=np.empty((3, n), np.float32), Pv=np.empty((9, n), np.float32), w=np.empty(n, np.float32))
        self._check(self._L.slamgpu_download_poses_at(self._h, idx.ctypes.data, n, _ptr(out["xv"]), _ptr(out["Pv"]),
                                                      _ptr(out["w"])))
        return out

    def download_slice(self, p0, n, landmarks=True):
        Nf = self.Nf
        out = dict(xv=np.empty((3, n), np.float32), Pv=np.empty((9, n), np.float32), w=np.empty(n, np.float32), Nf=Nf)
        if landmarks and Nf:
            out["lm_mean"] = np.empty((Nf, n, 2), np.float32)
            out["lm_cov"] = np.empty((Nf, n, 4), np.float32)
        self._check(self._L.slamgpu_download_slice(self._h, int(p0), int(n), _ptr(out["xv"]), _ptr(out["Pv"]),
                                                   _ptr(out["w"]), _ptr(out.get("lm_mean")), _ptr(out.get("lm_cov"))))
        return out

    def comm_init(self, rank, world, unique_id):
        buf = (C.c_char * 128).from_buffer_copy(bytes(unique_id))
        self._check(self._L.slamgpu_comm_init(self._h, rank, world, C.addressof(buf)))

    @staticmethod
    def comm_unique_id():
        L = _lib.load()
        buf = (C.c_char * 128)()
        rc = L.slamgpu_comm_unique_id(C.addressof(buf))
        if rc != 0:
            raise SlamGpuError(rc, L.slamgpu_last_error().decode())
        return bytes(buf)

    # -- state ----------------------------------------------------------------------
    def upload(self, xv=None, Pv=None, w=None, lm_mean=None, lm_cov=None, Nf=-1):
        xv, Pv, w, lm_mean, lm_cov = map(_f32, (xv, Pv, w, lm_mean, lm_cov))
        self._check(self._L.slamgpu_upload_particles(self._h, _ptr(xv), _ptr(Pv), _ptr(w), _ptr(lm_mean),
                                                     _ptr(lm_cov), int(Nf)))

    def upload_tiled(self, xv, Pv, w, lm_mean, lm_cov, Nf):
        """Replicate a tile of Pt = xv.shape[1] particles over all P (synthetic workloads)."""
        xv, Pv, w, lm_mean, lm_cov = map(_f32, (xv, Pv, w, lm_mean, lm_cov))
        self._check(self._L.slamgpu_upload_particles_tiled(self._h, xv.shape[1], _ptr(xv), _ptr(Pv), _ptr(w),
                                                           _ptr(lm_mean), _ptr(lm_cov), int(Nf)))

    def download(self, landmarks=True):
        P, Nf = self.P, self.Nf
        out = dict(xv=np.empty((3, P), np.float32), Pv=np.empty((9, P), np.float32), w=np.empty(P, np.float32), Nf=Nf)
        if landmarks:
            out["lm_mean"] = np.empty((Nf, P, 2), np.float32)
            out["lm_cov"] = np.empty((Nf, P, 4), np.float32)
        nf = C.c_int32()
        self._check(self._L.slamgpu_download_particles(
            self._h, _ptr(out["xv"]), _ptr(out["Pv"]), _ptr(out["w"]),
            _ptr(out["lm_mean"]) if landmarks and Nf else None, _ptr(out["lm_cov"]) if landmarks and Nf else None,
            C.byref(nf)))
        return out

    def stage_noise(self, gauss=None, unif=None):
        g = _f32(gauss)
        u = None if unif is None else np.ascontiguousarray(unif, np.float64)
        rows = 0 if g is None else g.shape[0]
        self._check(self._L.slamgpu_stage_noise(self._h, _ptr(g), rows, _ptr(u)))

    # -- the reference's entry points ---------------------------------------------------
    def predict(self, V, G, Q, dt, phi_true=0.0, gauss=None):
        """FastSLAM{1,2}::predict.  gauss: None or [2][P] standard normals (addPredictNoise)."""
        Q = _f32(Q).ravel()
        g = _f32(gauss)
        if g is not None:
            assert g.shape == (2, self.P)
        self._check(self._L.slamgpu_predict(self._h, V, G, _ptr(Q), dt, phi_true, _ptr(g)))

    def predict_raw(self, V, G, Q_ptr, dt, phi_true=0.0):
        """predict() without control noise on a raw host address for Q (no numpy marshalling)."""
        self._check(self._L.slamgpu_predict(self._h, V, G, Q_ptr, dt, phi_true, None))

    def _obs(self, zf, idf, zn):
        zf = _f32(np.asarray(zf, np.float32).reshape(-1, 2))
        zn = _f32(np.asarray(zn, np.float32).reshape(-1, 2))
        idf = np.ascontiguousarray(idf, np.int32)
        assert idf.shape[0] == zf.shape[0]
        return zf, idf, zn

    def update_particles(self, zf, idf, zn, R, gauss=None):
        zf, idf, zn = self._obs(zf, idf, zn)
        R = _f32(R).ravel()
        g = _f32(gauss)
        if g is not None:
            assert g.shape == (3, self.P)
        self._check(self._L.slamgpu_update_particles(self._h, _ptr(zf) if len(zf) else None,
                                                     _ptr(idf) if len(idf) else None, len(zf),
                                                     _ptr(zn) if len(zn) else None, len(zn), _ptr(R), _ptr(g)))

    def resample(self, unif=None, want_keep=False):
        """resampleParticles.  Returns (keep or None, nEff, did_resample)."""
        u = None if unif is None else np.ascontiguousarray(unif, np.float64)
        keep = np.empty(self.P, np.int32) if want_keep else None
        neff, did = C.c_float(), C.c_int32()
        self._check(self._L.slamgpu_resample(self._h, _ptr(u), _ptr(keep), C.byref(neff), C.byref(did)))
        return keep, neff.value, bool(did.value)

    def update(self, zf, idf, zn, R, gauss=None, unif=None):
        """FastSLAM{1,2}::update.  Returns (nEff, did_resample)."""
        zf, idf, zn = self._obs(zf, idf, zn)
        R = _f32(R).ravel()
        g = _f32(gauss)
        u = None if unif is None else np.ascontiguousarray(unif, np.float64)
        neff, did = C.c_float(), C.c_int32()
        self._check(self._L.slamgpu_update(self._h, _ptr(zf) if len(zf) else None, _ptr(idf) if len(idf) else None,
                                           len(zf), _ptr(zn) if len(zn) else None, len(zn), _ptr(R), _ptr(g), _ptr(u),
                                           C.byref(did), C.byref(neff)))
        return neff.value, bool(did.value)

    def update_raw(self, zf_ptr, idf_ptr, K, R_ptr, gauss_ptr, unif_ptr, unif_ints=False):
        """update() on raw host addresses (pinned buffers), no numpy marshalling: the benchmark's end-to-end call.
        unif_ints: unif_ptr holds the raw rand() ints (slamgpu_update_i32) instead of doubles."""
        neff, did = C.c_float(), C.c_int32()
        fn = self._L.slamgpu_update_i32 if unif_ints else self._L.slamgpu_update
        self._check(fn(self._h, zf_ptr, idf_ptr, K, None, 0, R_ptr, gauss_ptr, unif_ptr, C.byref(did), C.byref(neff)))
        return neff.value, bool(did.value)

    def update_i32(self, zf, idf, zn, R, gauss, rand_ints):
        """update() with the resample uniforms as the raw rand() values."""
        zf, idf, zn = self._obs(zf, idf, zn)
        R = _f32(R).ravel()
        g = _f32(gauss)
        r = np.ascontiguousarray(rand_ints, np.int32)
        neff, did = C.c_float(), C.c_int32()
        self._check(self._L.slamgpu_update_i32(self._h, _ptr(zf) if len(zf) else None, _ptr(idf) if len(idf) else None,
                                               len(zf), _ptr(zn) if len(zn) else None, len(zn), _ptr(R), _ptr(g),
                                               _ptr(r), C.byref(did), C.byref(neff)))
        return neff.value, bool(did.value)

    def compute_jacobians_batch(self, idf, R):
        idf = np.ascontiguousarray(idf, np.int32)
        R = _f32(R).ravel()
        out = np.empty((len(idf), self.P, 16), np.float32)
        self._check(self._L.slamgpu_compute_jacobians_batch(self._h, _ptr(idf), len(idf), _ptr(R), _ptr(out)))
        return out

    def estimate(self):
        x, y, t = C.c_double(), C.c_double(), C.c_double()
        self._check(self._L.slamgpu_estimate(self._h, C.byref(x), C.byref(y), C.byref(t)))
        return x.value, y.value, t.value

    def download_decimated(self, stride, landmarks=False):
        n = (self.P + stride - 1) // stride
        xv = np.empty((3, n), np.float32)
        lm = np.empty((self.Nf, n, 2), np.float32) if landmarks and self.Nf else None
        self._check(self._L.slamgpu_download_decimated(self._h, stride, _ptr(xv), _ptr(lm)))
        return xv, lm

    # -- timing -----------------------------------------------------------------------
    def timing_enable(self, on=True):
        self._check(self._L.slamgpu_timing_enable(self._h, int(on)))

    def timing_reset(self):
        self._check(self._L.slamgpu_timing_reset(self._h))

    def timing_read(self):
        out = {}
        for i, name in enumerate(T_NAMES):
            ms, ev, ln = C.c_double(), C.c_int64(), C.c_int64()
            self._check(self._L.slamgpu_timing_read(self._h, i, C.byref(ms), C.byref(ev), C.byref(ln)))
            out[name] = dict(ms=ms.value, events=ev.value, launches=ln.value)
        return out


def compute_jacobians(rec, n):
    """computeJacobians through the accelerator record of core.cpp:586-664 (one particle, n landmarks)."""
    L = _lib.load()
    rec = np.ascontiguousarray(rec, np.float32)
    assert rec.size >= 7 + 6 * n
    out = np.empty(16 * n, np.float32)
    rc = L.slamgpu_compute_jacobians(None, rec.ctypes.data, n, out.ctypes.data)
    if rc != 0:
        raise SlamGpuError(rc, L.slamgpu_last_error().decode())
    return out


NOALLOC_MAX_PARTICLES = 1 << 22


def estimate_noalloc(d_xv, d_w, n_particles, d_out, stream=0):
    """slamgpu_estimate_noalloc (include/slamgpu.h): the pose estimate on DEVICE pointers (ints), nothing allocated.
    d_xv [3][n] float32, d_w [n] float32, d_out 4 float64 — all device addresses on the current device; stream is a
    cudaStream_t handle (0 = default stream).  No counterpart in the reference."""
    L = _lib.load()
    rc = L.slamgpu_estimate_noalloc(int(d_xv), int(d_w), int(n_particles), int(d_out), int(stream) or None)
    if rc != 0:
        raise SlamGpuError(rc, L.slamgpu_last_error().decode())
