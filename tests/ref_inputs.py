"""Benchmark inputs for the CPU arms WITHOUT the product library: the sphere bytes come from committed fixtures
(tests/golden/*.bin, written by tests/golden/make_scene_bins.py from the seeded host generators and checked against them
in tests/test_abi.py) and the 176-byte camera uniform from the ORACLE's own restatement of SceneCamera::from
(src/camera.rs:293-351).  `bench.py --impl reference` therefore maps liboracle.so only.  Test infrastructure."""
import ctypes as C
import os

import numpy as np

import oracle_lib as ol

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SPHERE_DTYPE = np.dtype([("position", "<f4", 3), ("radius", "<f4"), ("color", "<f4", 4)])


class _Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class _CameraSettings(C.Structure):   # include/rt_abi.h CameraSettings == src/camera.rs:9-28
    _fields_ = [("field_of_view", C.c_float), ("samples_per_pixel", C.c_uint32), ("camera_has_moved", C.c_uint32),
                ("max_depth", C.c_uint32), ("vup", _Vec3), ("look_from", _Vec3), ("look_at", _Vec3),
                ("defocus_angle", C.c_float), ("focus_distance", C.c_float)]


class _SceneCamera(C.Structure):      # 176 bytes, src/camera.rs:256-291
    _fields_ = [("f", C.c_float * 44)]


class Scene:
    """Duck type of gpu_ray_tracing_b200.SphereCollection for oracle_lib.oracle_update."""

    def __init__(self, spheres, count):
        self.spheres, self.count = spheres, int(count)


def random_scene_1024():
    """The headline scene: rt_create_random_spheres(seed 1234, n 1024), as committed bytes."""
    raw = np.fromfile(os.path.join(GOLDEN, "random1024_seed1234.bin"), dtype=SPHERE_DTYPE)
    assert len(raw) == 1024
    return Scene(raw, 1024)


def camera(width, height, max_depth=1, spp=1, random_seed=0.5, moved=1):
    """CameraSettings::default() (src/camera.rs:30-46) with the given depth / sample cap, derived by the oracle."""
    s = _CameraSettings(20.0, spp, moved, max_depth, _Vec3(0, 1, 0), _Vec3(13, 2, 3), _Vec3(0, 0, 0), 0.6, 10.0)
    out = _SceneCamera()
    ol.oracle().oracle_camera_from_settings(C.byref(s), width, height, C.c_float(random_seed), C.byref(out))
    return out
