"""Host-side mirror of the reference's EKFSLAM method object (algorithms/ekfslam.h:14-56) over the C ABI
(include/slamgpu.h, slamgpu_ekf_*).  The joint state stays on the device; `sim` is EKFSLAM::sim."""
import ctypes as C

import numpy as np

from . import _lib
from .backend import SlamGpuError


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data


class EKFSLAM:
    """Same public knobs as the reference object: enableBatchUpdate, useHeading, wheelBase, gateReject, gateAugment,
    associationKnown, sigmaPhi (ekfslamwrapper.cpp:17-23)."""

    def __init__(self, nf_cap, wheel_base, sigma_phi, gate_reject=4.0, gate_augment=25.0, use_heading=True,
                 batch_update=True, association_known=False, m_cap=64, device=0):
        self._L = _lib.load()
        cfg = _lib.EkfConfig(nf_cap, m_cap, wheel_base, sigma_phi, gate_reject, gate_augment, int(use_heading),
                             int(batch_update), int(association_known), device)
        h = C.c_void_p()
        self._h = None
        self._check(self._L.slamgpu_ekf_create(C.byref(cfg), C.byref(h)))
        self._h = h

    def _check(self, rc):
        if rc != 0:
            raise SlamGpuError(rc, self._L.slamgpu_last_error().decode())

    def close(self):
        if self._h:
            self._L.slamgpu_ekf_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n(self):
        return self._L.slamgpu_ekf_size(self._h)

    @property
    def launches(self):
        return self._L.slamgpu_ekf_launch_count(self._h)

    def set(self, x, P):
        x, P = _f32(x), _f32(P)
        self._check(self._L.slamgpu_ekf_set_state(self._h, len(x), _ptr(x), _ptr(P)))

    def get(self):
        n = self.n
        x, P = np.empty(n, np.float32), np.empty((n, n), np.float32)
        self._check(self._L.slamgpu_ekf_get_state(self._h, _ptr(x), _ptr(P)))
        return x, P

    def pose(self):
        xv, Pvv = np.empty(3, np.float32), np.empty((3, 3), np.float32)
        self._check(self._L.slamgpu_ekf_get_pose(self._h, _ptr(xv), _ptr(Pvv)))
        return xv, Pvv

    def sim(self, V, G, Qe, dt, phi, observe, z=(), idz=None, table=None, Re=None, R=None):
        """EKFSLAM::sim (ekfslam.cpp:18-43).  Returns (idf, n_zf, n_zn) of this call's association."""
        z = _f32(np.asarray(z, np.float32).reshape(-1, 2))
        m = z.shape[0]
        idz = None if idz is None else np.ascontiguousarray(idz, np.int32)
        table = None if table is None else np.ascontiguousarray(table, np.int32)
        Qe = _f32(Qe).ravel()
        Re = None if Re is None else _f32(Re).ravel()
        R = None if R is None else _f32(R).ravel()
        idf = np.zeros(max(m, 1), np.int32)
        nzf, nzn = C.c_int32(), C.c_int32()
        self._check(self._L.slamgpu_ekf_sim(self._h, V, G, _ptr(Qe), dt, phi, int(observe), _ptr(z) if m else None,
                                            _ptr(idz), m, _ptr(table), 0 if table is None else len(table), _ptr(Re),
                                            _ptr(R), _ptr(idf), C.byref(nzf), C.byref(nzn)))
        return idf[:nzf.value].copy(), nzf.value, nzn.value
