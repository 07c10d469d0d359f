"""Loader for the CUDA shared library (librt_b200.so).

There is no CPU fallback: if the library is missing or older than its sources it is built with nvcc
(csrc/build.sh); if that fails, importing the bindings raises.  Entry points that need a device
return RT_ERR_NO_DEVICE on a machine without one and the Python layer turns that into RtError.
"""
import ctypes
import fcntl
import glob
import hashlib
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB: developer override, used by tools/ab_build.sh to time differently configured builds side by side
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")
_HASH_PATH = LIB_PATH + ".srchash"


def _source_hash() -> str:
    """Digest of every file the build reads: csrc/* (sources, headers, the build script) and include/*.  Content, not
    mtime: a copy of the tree (the GPU box gets one) keeps the library it was built with."""
    h = hashlib.sha256()
    files = sorted(glob.glob(os.path.join(_HERE, "csrc", "*")) + glob.glob(os.path.join(_HERE, "..", "include", "*")))
    for f in files:
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    if os.environ.get("RT_B200_LIB"):
        return False
    try:
        with open(_HASH_PATH) as f:
            return f.read().strip() != _source_hash()
    except OSError:
        return True


def build(force: bool = False) -> str:
    """Compile librt_b200.so for sm_100a (nvcc cross-compiles without a GPU) unless it is up to date with its sources."""
    with open(LIB_PATH + ".lock", "w") as lock:      # one builder at a time (torchrun starts N processes at once)
        fcntl.flock(lock, fcntl.LOCK_EX)
        if force or _stale():
            script = os.path.join(_HERE, "csrc", "build.sh")
            proc = subprocess.run(["sh", script], capture_output=True, text=True)
            if proc.returncode != 0:
                raise RuntimeError("building librt_b200.so failed:\n" + proc.stdout + proc.stderr)
            with open(_HASH_PATH, "w") as f:
                f.write(_source_hash())
    return LIB_PATH


def load() -> ctypes.CDLL:
    if _stale():
        build()
    return ctypes.CDLL(LIB_PATH)
