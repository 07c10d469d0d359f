"""ctypes access to the CPU oracle (oracle/liboracle*.so) and to the CPU emulation of the kernels
(tests/emu/libemu.so).  Test infrastructure only — never imported by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
EMU_DIR = os.path.join(ROOT, "tests", "emu")
_CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in src_list)


def build_oracle():
    deps = [os.path.join(ORACLE_DIR, "rt_oracle.cpp"), os.path.join(ROOT, "include", "rt_abi.h"),
            os.path.join(ROOT, "include", "rt_pinned_math.h")]
    if _newer(deps, os.path.join(ORACLE_DIR, "liboracle.so")) or _newer(deps, os.path.join(ORACLE_DIR, "liboracle_libm.so")):
        subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True, capture_output=True)


def build_emu():
    src = os.path.join(EMU_DIR, "emu_update.cpp")
    out = os.path.join(EMU_DIR, "libemu.so")
    deps = [src, os.path.join(ROOT, "gpu_ray_tracing_b200", "csrc", "rt_core.cuh"),
            os.path.join(ROOT, "gpu_ray_tracing_b200", "csrc", "rt_tree.h"),
            os.path.join(ROOT, "include", "rt_pinned_math.h"), os.path.join(ROOT, "include", "rt_abi.h")]
    if _newer(deps, out):
        subprocess.run([_CXX, "-std=c++17", "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared",
                        "-fvisibility=hidden", "-o", out, src], check=True, capture_output=True)


_P = C.c_void_p
_ORACLES = {}
_EMU = None


def oracle(libm: bool = False) -> C.CDLL:
    if libm not in _ORACLES:
        build_oracle()
        lib = C.CDLL(os.path.join(ORACLE_DIR, "liboracle_libm.so" if libm else "liboracle.so"))
        lib.oracle_hash.restype = C.c_uint32
        lib.oracle_hash.argtypes = [C.c_uint32]
        lib.oracle_random_float.restype = C.c_float
        lib.oracle_random_float.argtypes = [C.c_uint32]
        lib.oracle_sincos.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.oracle_pow5.restype = C.c_float
        lib.oracle_pow5.argtypes = [C.c_float]
        lib.oracle_random_unit_vector.argtypes = [C.c_uint32, _P]
        lib.oracle_sphere_hit.restype = C.c_int
        lib.oracle_sphere_hit.argtypes = [_P, _P, _P, C.c_float, C.c_float, _P]
        lib.oracle_sphere_list_hit.restype = C.c_int
        lib.oracle_sphere_list_hit.argtypes = [_P, C.c_uint32, _P, _P, C.c_float, C.c_float, _P, C.POINTER(C.c_int32)]
        lib.oracle_get_ray.argtypes = [_P, C.c_int32, C.c_int32, C.c_uint32, _P]
        lib.oracle_init.argtypes = [C.c_uint32, C.c_uint32, _P]
        lib.oracle_update.restype = C.c_uint64
        lib.oracle_update.argtypes = [_P, _P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, C.c_int]
        lib.oracle_quantize.argtypes = [_P, C.c_uint64, _P]
        lib.oracle_camera_from_settings.argtypes = [_P, C.c_uint32, C.c_uint32, C.c_float, _P]
        lib.oracle_max_threads.restype = C.c_int
        lib.oracle_uses_libm.restype = C.c_int
        _ORACLES[libm] = lib
    return _ORACLES[libm]


def emu() -> C.CDLL:
    global _EMU
    if _EMU is None:
        build_emu()
        lib = C.CDLL(os.path.join(EMU_DIR, "libemu.so"))
        lib.emu_update.argtypes = [_P, _P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, _P, C.c_int]
        _EMU = lib
    return _EMU


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle_update(camera, scene, width, height, accum=None, row_begin=0, row_count=None, threads=0, libm=False):
    """One WGSL `update` over the window on the CPU.  Returns (accum, rays)."""
    if row_count is None:
        row_count = height - row_begin
    if accum is None:
        accum = np.zeros((row_count, width, 4), dtype=np.float32)
    assert accum.dtype == np.float32 and accum.flags.c_contiguous and accum.size == row_count * width * 4
    rays = oracle(libm).oracle_update(C.byref(camera), _p(scene.spheres), scene.count, width, height, row_begin,
                                      row_count, _p(accum), threads)
    return accum, int(rays)


EMU_STATS = {}


def emu_update(camera, scene, width, height, accum=None, row_begin=0, row_count=None, variant=0):
    """The kernels' control flow + arithmetic on the CPU.  Returns (accum, rays, exact_tests, violations);
    EMU_STATS holds the primary-pass counters of the last call.  variant as rt_set_kernel_variant (0/4 use the
    common-origin primary pass when it applies, 2 the general broad phase from bounce 0)."""
    if row_count is None:
        row_count = height - row_begin
    if accum is None:
        accum = np.zeros((row_count, width, 4), dtype=np.float32)
    stats = np.zeros(13, dtype=np.uint64)
    emu().emu_update(C.byref(camera), _p(scene.spheres), scene.count, width, height, row_begin, row_count,
                     _p(accum), _p(stats), variant)
    EMU_STATS.update(primary_rays=int(stats[3]), untraced=int(stats[4]), tile_pairs=int(stats[5]), tile_dropped=int(stats[6]),
                     cluster_rays=int(stats[7]), cluster_tests=int(stats[8]), member_tests=int(stats[9]), retraced=int(stats[10]), tile_steps=int(stats[11]), beaten=int(stats[12]))
    return accum, int(stats[0]), int(stats[1]), int(stats[2])


def quantize(accum):
    out = np.empty(accum.shape[:-1] + (4,), dtype=np.uint8)
    a = np.ascontiguousarray(accum, dtype=np.float32)
    oracle().oracle_quantize(_p(a), a.size // 4, _p(out))
    return out


def compare_images(a, b):
    """Parity report between two rgba32f accumulations: bit-identical flag, max |delta| in float,
    and the count of pixels whose RGBA8 quantisation differs by more than 1 in any channel."""
    same_bits = bool(np.array_equal(a.view(np.uint32), b.view(np.uint32)))
    fa, fb = a[..., :3].astype(np.float64), b[..., :3].astype(np.float64)
    max_abs = float(np.nanmax(np.abs(fa - fb))) if a.size else 0.0
    qa, qb = quantize(a).astype(np.int16), quantize(b).astype(np.int16)
    dq = np.abs(qa[..., :3] - qb[..., :3]).max(axis=-1)
    return {"bit_identical": same_bits, "max_abs": max_abs,
            "pixels_over_1_255": int((dq > 1).sum()), "pixels_differing": int((dq > 0).sum()),
            "pixels": int(dq.size)}


def nan_aware_equal(a, b):
    return np.array_equal(a.view(np.uint32)[~np.isnan(a)], b.view(np.uint32)[~np.isnan(a)]) and \
        np.array_equal(np.isnan(a), np.isnan(b))
