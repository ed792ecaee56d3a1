"""Builds and loads tools/probes/libaz_probes.so: the measurement-only kernels (az_peak_int32, az_peak_mma)."""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libaz_probes.so")
SOURCES = ["peak_kernels.cu", "mma_probe.cu", "mma2_probe.cu", "probe_common.cu"]
_lib = None


def build() -> str:
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
           "-shared", "-o", LIB] + [os.path.join(HERE, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    return LIB


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        srcs = [os.path.join(HERE, s) for s in SOURCES]
        if not os.path.isfile(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
            build()
        L = ctypes.CDLL(LIB)
        c_int, c_i64, c_void_p = ctypes.c_int, ctypes.c_int64, ctypes.c_void_p
        L.az_peak_int32.restype = c_int
        L.az_peak_int32.argtypes = [c_int, c_int, c_void_p, c_i64, ctypes.POINTER(ctypes.c_int64), c_void_p]
        L.az_peak_mma.restype = c_int
        L.az_peak_mma.argtypes = [c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]
        L.az_probe_mma2.restype = c_int
        L.az_probe_mma2.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]
        _lib = L
    return _lib
