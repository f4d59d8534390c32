"""Builds slam_b200/libslamgpu.so (the C-ABI CUDA library of include/slamgpu.h) in-tree with nvcc for sm_100a.

nvcc cross-compiles without a GPU; the built file travels to the GPU box with the repo snapshot."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libslamgpu.so")
SOURCES = ["capi.cu", "comm.cu", "ekf.cu", "noalloc.cu"]
HEADERS = ["slam_math.cuh", "slam_rng.cuh", "xchg.cuh", "kernels.cuh", "ekf_kernels.cuh", "types.cuh", "handle.h", os.path.join("..", "..", "include", "slamgpu.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-shared"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    """Compile the library if any source is newer than it.  Returns the library path."""
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + \
          [os.path.join(CSRC, f) for f in SOURCES] + ["-ldl"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
