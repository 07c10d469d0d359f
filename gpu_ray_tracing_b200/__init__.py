"""B200-native drop-in for the per-pixel ray-trace pass of Sur091/GPU-Ray-Tracing.

The product is the CUDA shared library behind include/rt_abi.h; this package is the thin host
layer that mirrors the reference's names for tests, benchmarks and headless use.
"""
from .api import (  # noqa: F401
    CameraSettings, ComputeShaderNode, Key, camera_movement_system, camera_script, ComputeShaderState, GpuSphere, MultiRenderer, Renderer, RtError, RtVec3,
    SceneCamera, SphereCollection, SPHERE_DTYPE, RT_PRECISION_EXACT, RT_PRECISION_FAST_SHADING, RT_PRECISION_FAST_ALL, create_default_spheres, create_grid_spheres,
    create_random_spheres, device_alloc, device_count, device_free, ipc_close, ipc_export, ipc_open, lib, orbit_camera, render, render_frames, update_static, RT_STATIC_MAX_SPHERES, shard_bands, shard_rows, write_ppm, write_png, read_png, diff_image,
)
from ._lib import LIB_PATH, build  # noqa: F401
