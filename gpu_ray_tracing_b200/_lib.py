"""Loader for the CUDA shared library (librt_b200.so).

There is no CPU fallback: if the library is missing it is built with nvcc (csrc/build.sh); if that
fails, importing the bindings raises.  Entry points that need a device return RT_ERR_NO_DEVICE on
a machine without one and the Python layer turns that into RtError.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB: developer override, used by tools/ab_build.sh to time differently configured builds side by side
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")
_SOURCES = ["rt_kernels.cu", "rt_kernels.cuh", "rt_core.cuh", "rt_api.cu", "rt_multi.cu", "rt_host.cpp",
            "rt_internal.h", "rt_internal_host.h", "build.sh"]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(_HERE, "csrc", s) for s in _SOURCES]
    deps += [os.path.join(_HERE, "..", "include", h) for h in ("rt_abi.h", "rt_pinned_math.h")]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False) -> str:
    """Compile librt_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    if force or _stale():
        script = os.path.join(_HERE, "csrc", "build.sh")
        proc = subprocess.run(["sh", script], capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("building librt_b200.so failed:\n" + proc.stdout + proc.stderr)
    return LIB_PATH


def load() -> ctypes.CDLL:
    if not os.path.exists(LIB_PATH):
        build()
    return ctypes.CDLL(LIB_PATH)
