"""TEST INFRASTRUCTURE ONLY — ctypes loaders for the EKF-SLAM checkers.

* ``EkfOracle`` : oracle/liboracle.so, the plain-C restatement (oracle/ekf_oracle.c)
* ``EkfRef``    : oracle/_ref/libslamref.so, the UNMODIFIED reference EKFSLAM through oracle/ref_shim_ekf.cpp

Both expose ``sim(V, G, Qe, dt, phi, observe, z, idz, table, Re, R) -> (idf, n_zf, n_zn)`` (EKFSLAM::sim,
ekfslam.cpp:18-43) and ``get() / set(x, P)`` with P row-major [n][n].  Only tests/ import this.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def _f(a):
    return np.ascontiguousarray(a, np.float32)


def _i(a):
    return np.ascontiguousarray(a, np.int32)


class _Cfg(C.Structure):
    _fields_ = [("wheel_base", C.c_float), ("sigma_phi", C.c_float), ("gate_reject", C.c_float),
                ("gate_augment", C.c_float), ("use_heading", C.c_int), ("batch_update", C.c_int),
                ("association_known", C.c_int)]


def _obs(z, idz):
    z = _f(np.asarray(z, np.float32).reshape(-1, 2))
    m = z.shape[0]
    idz = _i(idz if idz is not None else np.zeros(m, np.int32))
    return (z if m else np.zeros((1, 2), np.float32)), (idz if m else np.zeros(1, np.int32)), m


class EkfOracle:
    def __init__(self, wheel_base, sigma_phi, gate_reject, gate_augment, use_heading=True, batch_update=True,
                 association_known=False, cap=512):
        L = self.L = C.CDLL(os.path.join(HERE, "liboracle.so"))
        self.cfg = _Cfg(wheel_base, sigma_phi, gate_reject, gate_augment, int(use_heading), int(batch_update),
                        int(association_known))
        L.so_ekf_sim.restype = C.c_int
        L.so_ekf_sim.argtypes = [C.POINTER(_Cfg), C.c_int, C.c_int, f32p, f32p, C.c_float, C.c_float, f32p, C.c_float,
                                 C.c_float, C.c_int, f32p, i32p, C.c_int, C.c_void_p, f32p, f32p, i32p,
                                 C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.so_ekf_compute_association.argtypes = [C.c_int, f32p, f32p, f32p, f32p, C.c_int, C.POINTER(C.c_float),
                                                 C.POINTER(C.c_float)]
        self.ld = cap
        self.n = 3
        self.x = np.zeros(cap, np.float32)
        self.P = np.zeros((cap, cap), np.float32)

    def set(self, x, P):
        n = len(x)
        self.n = n
        self.x[:] = 0
        self.P[:] = 0
        self.x[:n] = x
        self.P[:n, :n] = P

    def get(self):
        return self.x[:self.n].copy(), self.P[:self.n, :self.n].copy()

    def sim(self, V, G, Qe, dt, phi, observe, z=(), idz=None, table=None, Re=None, R=None):
        z, idz, m = _obs(z, idz)
        idf = np.zeros(max(m, 1), np.int32)
        nzf, nzn = C.c_int(), C.c_int()
        tab = None if table is None else table.ctypes.data_as(C.c_void_p)
        self.n = self.L.so_ekf_sim(C.byref(self.cfg), self.n, self.ld, self.x, self.P, V, G, _f(Qe).ravel(), dt, phi,
                                   int(observe), z, idz, m, tab, _f(Re).ravel(), _f(R).ravel(), idf, C.byref(nzf),
                                   C.byref(nzn))
        return idf[:nzf.value].copy(), nzf.value, nzn.value

    def association(self, z, R, idf):
        nis, nd = C.c_float(), C.c_float()
        self.L.so_ekf_compute_association(self.ld, self.x, self.P, _f(z), _f(R).ravel(), idf, C.byref(nis), C.byref(nd))
        return nis.value, nd.value


class EkfRef:
    @staticmethod
    def available():
        return os.path.exists(os.path.join(HERE, "_ref", "libslamref.so"))

    def __init__(self, wheel_base, sigma_phi, gate_reject, gate_augment, use_heading=True, batch_update=True,
                 association_known=False):
        L = self.L = C.CDLL(os.path.join(HERE, "_ref", "libslamref.so"))
        L.slamref_ekf_create.restype = C.c_void_p
        L.slamref_ekf_create.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int]
        L.slamref_ekf_destroy.argtypes = [C.c_void_p]
        L.slamref_ekf_size.argtypes = [C.c_void_p]
        L.slamref_ekf_set.argtypes = [C.c_void_p, C.c_int, f32p, f32p]
        L.slamref_ekf_get.argtypes = [C.c_void_p, f32p, f32p]
        L.slamref_ekf_sim.argtypes = [C.c_void_p, C.c_float, C.c_float, f32p, C.c_float, C.c_float, C.c_int, f32p, i32p,
                                      C.c_int, C.c_void_p, C.c_int, f32p, f32p, i32p, C.POINTER(C.c_int),
                                      C.POINTER(C.c_int)]
        L.slamref_ekf_association.argtypes = [C.c_void_p, f32p, f32p, C.c_int, C.POINTER(C.c_float),
                                              C.POINTER(C.c_float)]
        self.h = L.slamref_ekf_create(wheel_base, sigma_phi, gate_reject, gate_augment, int(use_heading),
                                      int(batch_update), int(association_known))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.slamref_ekf_destroy(self.h)
            self.h = None

    @property
    def n(self):
        return self.L.slamref_ekf_size(self.h)

    def set(self, x, P):
        self.L.slamref_ekf_set(self.h, len(x), _f(x), _f(P))

    def get(self):
        n = self.n
        x, P = np.zeros(n, np.float32), np.zeros((n, n), np.float32)
        self.L.slamref_ekf_get(self.h, x, P)
        return x, P

    def sim(self, V, G, Qe, dt, phi, observe, z=(), idz=None, table=None, Re=None, R=None):
        z, idz, m = _obs(z, idz)
        idf = np.zeros(max(m, 1), np.int32)
        nzf, nzn = C.c_int(), C.c_int()
        tab = None if table is None else table.ctypes.data_as(C.c_void_p)
        self.L.slamref_ekf_sim(self.h, V, G, _f(Qe).ravel(), dt, phi, int(observe), z, idz, m, tab,
                               0 if table is None else len(table), _f(Re).ravel(), _f(R).ravel(), idf, C.byref(nzf),
                               C.byref(nzn))
        return idf[:nzf.value].copy(), nzf.value, nzn.value

    def association(self, z, R, idf):
        nis, nd = C.c_float(), C.c_float()
        self.L.slamref_ekf_association(self.h, _f(z), _f(R).ravel(), idf, C.byref(nis), C.byref(nd))
        return nis.value, nd.value
