"""ctypes declarations for include/slamgpu.h.  The library is the product: if it is missing this raises, there is no
Python or CPU stand-in."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libslamgpu.so")

# Every symbol include/slamgpu.h declares (checked by tests/test_abi.py).
EXPORTS = [
    "slamgpu_last_error", "slamgpu_version", "slamgpu_create", "slamgpu_destroy", "slamgpu_set_stream",
    "slamgpu_get_stream", "slamgpu_synchronize", "slamgpu_host_alloc", "slamgpu_host_free",
    "slamgpu_upload_particles", "slamgpu_upload_particles_tiled", "slamgpu_download_particles", "slamgpu_landmark_count", "slamgpu_predict",
    "slamgpu_update", "slamgpu_update_particles", "slamgpu_resample", "slamgpu_stage_noise",
    "slamgpu_compute_jacobians", "slamgpu_compute_jacobians_batch", "slamgpu_estimate", "slamgpu_download_decimated",
    "slamgpu_timing_enable", "slamgpu_timing_read", "slamgpu_timing_reset", "slamgpu_launch_count", "slamgpu_stat",
    "slamgpu_comm_unique_id", "slamgpu_comm_init", "slamgpu_update_i32", "slamgpu_seed", "slamgpu_download_slice",
    "slamgpu_debug_capture", "slamgpu_debug_read", "slamgpu_download_poses_at", "slamgpu_estimate_noalloc",
    "slamgpu_ekf_create", "slamgpu_ekf_destroy", "slamgpu_ekf_sim", "slamgpu_ekf_size", "slamgpu_ekf_set_state",
    "slamgpu_ekf_get_state", "slamgpu_ekf_get_pose", "slamgpu_ekf_launch_count",
]


class EkfConfig(C.Structure):
    _fields_ = [("nf_cap", C.c_int32), ("m_cap", C.c_int32), ("wheel_base", C.c_float), ("sigma_phi", C.c_float),
                ("gate_reject", C.c_float), ("gate_augment", C.c_float), ("use_heading", C.c_int32),
                ("batch_update", C.c_int32), ("association_known", C.c_int32), ("device", C.c_int32),
                ("reserved", C.c_int32 * 4)]


class Config(C.Structure):
    _fields_ = [("n_particles", C.c_int64), ("nf_cap", C.c_int32), ("method", C.c_int32), ("wheel_base", C.c_float),
                ("sigma_phi", C.c_float), ("n_effective", C.c_int32), ("use_heading", C.c_int32),
                ("do_resample", C.c_int32), ("scan_mode", C.c_int32), ("device", C.c_int32), ("spare_rows", C.c_int32),
                ("math_mode", C.c_int32), ("log_epochs", C.c_int32), ("keep_nan_weights", C.c_int32),
                ("reserved", C.c_int32 * 1)]


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("slam_b200/libslamgpu.so is not built (python -m slam_b200.build); "
                           "slam_b200 has no fallback path")
    L = C.CDLL(LIB_PATH)
    vp, fp, ip, dp = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p  # raw addresses; wrappers pass numpy .ctypes.data
    L.slamgpu_ekf_create.argtypes = [C.POINTER(EkfConfig), C.POINTER(vp)]
    L.slamgpu_ekf_destroy.argtypes = [vp]
    L.slamgpu_ekf_destroy.restype = None
    L.slamgpu_ekf_sim.argtypes = [vp, C.c_float, C.c_float, fp, C.c_float, C.c_float, C.c_int32, fp, ip, C.c_int32, ip,
                                  C.c_int32, fp, fp, ip, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.slamgpu_ekf_size.argtypes = [vp]
    L.slamgpu_ekf_set_state.argtypes = [vp, C.c_int32, fp, fp]
    L.slamgpu_ekf_get_state.argtypes = [vp, fp, fp]
    L.slamgpu_ekf_get_pose.argtypes = [vp, fp, fp]
    L.slamgpu_ekf_launch_count.argtypes = [vp]
    L.slamgpu_ekf_launch_count.restype = C.c_int64
    L.slamgpu_last_error.restype = C.c_char_p
    L.slamgpu_version.restype = C.c_char_p
    L.slamgpu_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.slamgpu_destroy.argtypes = [vp]
    L.slamgpu_destroy.restype = None
    L.slamgpu_set_stream.argtypes = [vp, vp]
    L.slamgpu_get_stream.argtypes = [vp]
    L.slamgpu_get_stream.restype = vp
    L.slamgpu_synchronize.argtypes = [vp]
    L.slamgpu_host_alloc.argtypes = [C.c_uint64]
    L.slamgpu_host_alloc.restype = vp
    L.slamgpu_host_free.argtypes = [vp]
    L.slamgpu_host_free.restype = None
    L.slamgpu_upload_particles.argtypes = [vp, fp, fp, fp, fp, fp, C.c_int32]
    L.slamgpu_upload_particles_tiled.argtypes = [vp, C.c_int64, fp, fp, fp, fp, fp, C.c_int32]
    L.slamgpu_download_particles.argtypes = [vp, fp, fp, fp, fp, fp, C.POINTER(C.c_int32)]
    L.slamgpu_landmark_count.argtypes = [vp]
    L.slamgpu_predict.argtypes = [vp, C.c_float, C.c_float, fp, C.c_float, C.c_float, fp]
    L.slamgpu_update.argtypes = [vp, fp, ip, C.c_int32, fp, C.c_int32, fp, fp, dp, C.POINTER(C.c_int32),
                                 C.POINTER(C.c_float)]
    L.slamgpu_update_i32.argtypes = [vp, fp, ip, C.c_int32, fp, C.c_int32, fp, fp, ip, C.POINTER(C.c_int32),
                                     C.POINTER(C.c_float)]
    L.slamgpu_seed.argtypes = [vp, C.c_uint64, C.c_int32]
    L.slamgpu_download_slice.argtypes = [vp, C.c_int64, C.c_int64, fp, fp, fp, fp, fp]
    L.slamgpu_download_poses_at.argtypes = [vp, ip, C.c_int64, fp, fp, fp]
    L.slamgpu_debug_capture.argtypes = [vp, C.c_int32]
    L.slamgpu_debug_read.argtypes = [vp, C.c_int32, vp]
    L.slamgpu_update_particles.argtypes = [vp, fp, ip, C.c_int32, fp, C.c_int32, fp, fp]
    L.slamgpu_resample.argtypes = [vp, dp, ip, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
    L.slamgpu_stage_noise.argtypes = [vp, fp, C.c_int32, dp]
    L.slamgpu_compute_jacobians.argtypes = [vp, fp, C.c_int32, fp]
    L.slamgpu_compute_jacobians_batch.argtypes = [vp, ip, C.c_int32, fp, fp]
    L.slamgpu_estimate.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.slamgpu_estimate_noalloc.argtypes = [vp, vp, C.c_int64, vp, vp]
    L.slamgpu_download_decimated.argtypes = [vp, C.c_int64, fp, fp]
    L.slamgpu_timing_enable.argtypes = [vp, C.c_int32]
    L.slamgpu_timing_read.argtypes = [vp, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int64)]
    L.slamgpu_timing_reset.argtypes = [vp]
    L.slamgpu_launch_count.argtypes = [vp]
    L.slamgpu_launch_count.restype = C.c_int64
    L.slamgpu_stat.argtypes = [vp, C.c_int32]
    L.slamgpu_stat.restype = C.c_int64
    L.slamgpu_comm_unique_id.argtypes = [vp]
    L.slamgpu_comm_init.argtypes = [vp, C.c_int32, C.c_int32, vp]
    _lib = L
    return L
