"""ctypes bindings for include/rt_abi.h plus a thin host layer that mirrors the reference's names.

Reference counterparts (paths relative to the reference repo):
  CameraSettings / SceneCamera ............ src/camera.rs:9-46, 256-351
  GpuSphere / SphereCollection ............ src/scene/sphere.rs:14-42
  create_default_spheres .................. src/scene/sphere.rs:45-153 (seeded here)
  ComputeShaderNode (Loading/Init/Update) . src/lib.rs:326-422
  render(scene, camera, w, h) ............. new headless entry (rt_render)

Everything that computes goes through the C ABI; numpy is used only to hold bytes.
"""
import ctypes as C
import enum

import numpy as np

from . import _lib

# ------------------------------------------------------------------------------------------------
# layouts
# ------------------------------------------------------------------------------------------------


class RtVec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    def __init__(self, x=0.0, y=0.0, z=0.0):
        super().__init__(x, y, z)

    def tuple(self):
        return (self.x, self.y, self.z)


class SceneCamera(C.Structure):
    """176-byte uniform, src/camera.rs:256-291."""
    _fields_ = [
        ("center", RtVec3), ("viewport_height", C.c_float),
        ("viewport_upper_left", RtVec3), ("viewport_width", C.c_float),
        ("pixel_delta_u", RtVec3), ("defocus_angle", C.c_float),
        ("pixel_delta_v", RtVec3), ("aspect_ratio", C.c_float),
        ("defocus_disk_u", RtVec3), ("_padding0", C.c_float),
        ("viewport_u", RtVec3), ("_padding1", C.c_float),
        ("defocus_disk_v", RtVec3), ("max_depth", C.c_float),
        ("look_from", RtVec3), ("samples_per_pixel", C.c_float),
        ("look_at", RtVec3), ("camera_has_moved", C.c_float),
        ("vup", RtVec3), ("random_seed", C.c_float),
        ("viewport_v", RtVec3), ("defocus_radius", C.c_float),
    ]

    def copy(self):
        out = SceneCamera()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(SceneCamera))
        return out

    def to_bytes(self) -> bytes:
        return bytes(memoryview(self))

    @classmethod
    def from_bytes(cls, b: bytes) -> "SceneCamera":
        assert len(b) == 176
        return cls.from_buffer_copy(b)

    @classmethod
    def from_settings(cls, settings: "CameraSettings", width: int, height: int, random_seed: float) -> "SceneCamera":
        """SceneCamera::from(&CameraSettings), src/camera.rs:293-351, for a width x height image."""
        out = cls()
        _check(lib().rt_camera_from_settings(C.byref(settings), width, height, C.c_float(random_seed), C.byref(out)))
        return out


class GpuSphere(C.Structure):
    """32-byte storage element, src/scene/sphere.rs:20-26."""
    _fields_ = [("position", RtVec3), ("radius", C.c_float), ("color", C.c_float * 4)]


class CameraSettings(C.Structure):
    """src/camera.rs:9-28; defaults from camera.rs:30-46."""
    _fields_ = [
        ("field_of_view", C.c_float), ("samples_per_pixel", C.c_uint32), ("camera_has_moved", C.c_uint32),
        ("max_depth", C.c_uint32), ("vup", RtVec3), ("look_from", RtVec3), ("look_at", RtVec3),
        ("defocus_angle", C.c_float), ("focus_distance", C.c_float),
    ]

    @classmethod
    def default(cls) -> "CameraSettings":
        s = cls()
        lib().rt_camera_settings_default(C.byref(s))
        return s


SPHERE_DTYPE = np.dtype([("position", "<f4", 3), ("radius", "<f4"), ("color", "<f4", 4)])
assert SPHERE_DTYPE.itemsize == 32 and C.sizeof(GpuSphere) == 32 and C.sizeof(SceneCamera) == 176

RT_OK = 0
RT_PRECISION_EXACT, RT_PRECISION_FAST_SHADING, RT_PRECISION_FAST_ALL = 0, 1, 2
RT_ERR_INVALID_ARG, RT_ERR_CUDA, RT_ERR_NO_DEVICE, RT_ERR_NO_SCENE, RT_ERR_OUT_OF_MEMORY, RT_ERR_STATE = -1, -2, -3, -4, -5, -6


class RtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"rt error {code}: {message}")
        self.code = code


_LIB = None

# name -> (restype, argtypes); every symbol include/rt_abi.h declares
_P = C.c_void_p
_SIGNATURES = {
    "rt_last_error": (C.c_char_p, []),
    "rt_version": (C.c_char_p, []),
    "rt_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "rt_create": (C.c_int, [C.c_int, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "rt_create_window": (C.c_int, [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "rt_create_bands": (C.c_int, [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "rt_shard_bands": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]),
    "rt_bands": (C.c_int, [_P, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "rt_destroy": (C.c_int, [_P]),
    "rt_set_scene": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32]),
    "rt_init": (C.c_int, [_P]),
    "rt_update": (C.c_int, [_P, C.POINTER(SceneCamera)]),
    "rt_update_n": (C.c_int, [_P, C.POINTER(SceneCamera), C.c_uint32]),
    "rt_synchronize": (C.c_int, [_P]),
    "rt_read_accum": (C.c_int, [_P, _P]),
    "rt_write_accum": (C.c_int, [_P, _P]),
    "rt_read_rgba8": (C.c_int, [_P, _P]),
    "rt_read_rgba8_async": (C.c_int, [_P, _P]),
    "rt_wavefront_counts": (C.c_int, [_P, _P, C.c_uint32, _P]),
    "rt_tree_info": (C.c_int, [_P, _P, _P, _P]),
    "rt_read_wait": (C.c_int, [_P]),
    "rt_set_rgba8_output": (C.c_int, [_P, C.c_int]),
    "rt_set_rgba8_target": (C.c_int, [_P, _P]),
    "rt_stream_signal": (C.c_int, [_P, _P, C.c_uint32]),
    "rt_stream_wait": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32, C.c_uint32]),
    "rt_stream_wait_timed_out": (C.c_int, [_P, C.POINTER(C.c_uint32)]),
    "rt_send_rgba8_async": (C.c_int, [_P, _P]),
    "rt_copy_signal": (C.c_int, [_P, _P, C.c_uint32]),
    "rt_copy_wait": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32, C.c_uint32]),
    "rt_set_precision": (C.c_int, [_P, C.c_int]),
    "rt_tree_build_ms": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "rt_device_alloc": (C.c_int, [C.c_int, C.c_uint64, C.POINTER(_P)]),
    "rt_device_free": (C.c_int, [C.c_int, _P]),
    "rt_ipc_export": (C.c_int, [C.c_int, _P, _P]),
    "rt_ipc_open": (C.c_int, [C.c_int, _P, C.POINTER(_P)]),
    "rt_ipc_close": (C.c_int, [C.c_int, _P]),
    "rt_rays_traced": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "rt_exact_tests": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "rt_reset_counters": (C.c_int, [_P]),
    "rt_timer_begin": (C.c_int, [_P]),
    "rt_timer_end": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "rt_kernel_launches": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "rt_accum_device_ptr": (C.c_int, [_P, C.POINTER(_P)]),
    "rt_rgba8_device_ptr": (C.c_int, [_P, C.POINTER(_P)]),
    "rt_quantize": (C.c_int, [_P]),
    "rt_quantize_into": (C.c_int, [_P, _P]),
    "rt_stream": (C.c_int, [_P, C.POINTER(_P)]),
    "rt_window": (C.c_int, [_P] + [C.POINTER(C.c_uint32)] * 4),
    "rt_set_kernel_variant": (C.c_int, [_P, C.c_int]),
    "rt_update_static": (C.c_int, [C.c_int, _P, C.c_uint32, C.POINTER(SceneCamera), C.c_uint32, C.c_uint32, _P, _P]),
    "rt_render": (C.c_int, [C.c_int, _P, C.c_uint32, C.POINTER(SceneCamera), C.c_uint32, C.c_uint32, C.c_uint32, _P]),
    "rt_render_frames": (C.c_int, [C.c_int, _P, C.c_uint32, C.POINTER(SceneCamera), C.c_uint32, C.c_uint32, C.c_uint32, _P]),
    "rt_shard_rows": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "rt_multi_create": (C.c_int, [C.POINTER(C.c_int), C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "rt_multi_create_bands": (C.c_int, [C.POINTER(C.c_int), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    "rt_multi_destroy": (C.c_int, [_P]),
    "rt_multi_set_scene": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32]),
    "rt_multi_init": (C.c_int, [_P]),
    "rt_multi_update": (C.c_int, [_P, C.POINTER(SceneCamera)]),
    "rt_multi_gather_rgba8": (C.c_int, [_P, _P]),
    "rt_multi_gather_accum": (C.c_int, [_P, _P]),
    "rt_multi_rays_traced": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "rt_multi_shard": (C.c_int, [_P, C.c_uint32, C.POINTER(_P)]),
    "rt_camera_settings_default": (None, [C.POINTER(CameraSettings)]),
    "rt_camera_from_settings": (C.c_int, [C.POINTER(CameraSettings), C.c_uint32, C.c_uint32, C.c_float, C.POINTER(SceneCamera)]),
    "rt_create_default_spheres": (C.c_int, [C.c_uint64, _P, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "rt_create_random_spheres": (C.c_int, [C.c_uint64, C.c_uint32, _P]),
    "rt_create_grid_spheres": (C.c_int, [C.c_uint64, C.c_uint32, _P]),
    "rt_orbit_camera": (C.c_int, [C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.POINTER(SceneCamera)]),
    "rt_camera_step": (C.c_int, [C.POINTER(CameraSettings), C.c_uint32, C.c_float]),
    "rt_camera_script": (C.c_int, [C.POINTER(CameraSettings), C.POINTER(C.c_uint32), C.c_uint32, C.c_float, C.POINTER(C.c_float),
                                   C.c_uint32, C.c_uint32, C.POINTER(SceneCamera)]),
}


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        handle = _lib.load()
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError here == the library does not export the ABI
            fn.restype = restype
            fn.argtypes = argtypes
        _LIB = handle
    return _LIB


def _check(code: int) -> None:
    if code != RT_OK:
        raise RtError(code, lib().rt_last_error().decode("utf-8", "replace"))


def device_count() -> int:
    n = C.c_int(0)
    code = lib().rt_device_count(C.byref(n))
    return n.value if code == RT_OK else 0


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


# ------------------------------------------------------------------------------------------------
# scenes (src/scene/sphere.rs)
# ------------------------------------------------------------------------------------------------
class SphereCollection:
    """`SphereCollection { spheres, count }` (sphere.rs:29-33): a padded buffer and the live count."""

    def __init__(self, spheres: np.ndarray, count: int):
        spheres = np.ascontiguousarray(spheres, dtype=SPHERE_DTYPE)
        if count > len(spheres):
            raise ValueError("count exceeds buffer length")
        self.spheres = spheres
        self.count = int(count)

    def to_bytes(self) -> bytes:
        return self.spheres.tobytes()

    @classmethod
    def from_bytes(cls, b: bytes, count: int) -> "SphereCollection":
        return cls(np.frombuffer(b, dtype=SPHERE_DTYPE).copy(), count)


def create_default_spheres(seed: int = 0) -> SphereCollection:
    """sphere.rs:45-153 with a seeded generator in place of rand::random."""
    buf = np.zeros(256, dtype=SPHERE_DTYPE)
    count, length = C.c_uint32(0), C.c_uint32(0)
    _check(lib().rt_create_default_spheres(seed, _ptr(buf), len(buf), C.byref(count), C.byref(length)))
    return SphereCollection(buf[: length.value].copy(), count.value)


def create_random_spheres(n: int, seed: int = 1234) -> SphereCollection:
    buf = np.zeros(n, dtype=SPHERE_DTYPE)
    _check(lib().rt_create_random_spheres(seed, n, _ptr(buf)))
    return SphereCollection(buf, n)


def create_grid_spheres(n: int, seed: int = 4321) -> SphereCollection:
    buf = np.zeros(n, dtype=SPHERE_DTYPE)
    _check(lib().rt_create_grid_spheres(seed, n, _ptr(buf)))
    return SphereCollection(buf, n)


def device_alloc(device: int, nbytes: int) -> int:
    p = _P()
    _check(lib().rt_device_alloc(device, nbytes, C.byref(p)))
    return p.value


def device_free(device: int, ptr: int) -> None:
    _check(lib().rt_device_free(device, ptr))


def ipc_export(device: int, ptr: int) -> bytes:
    h = (C.c_uint8 * 64)()
    _check(lib().rt_ipc_export(device, ptr, h))
    return bytes(h)


def ipc_open(device: int, handle: bytes) -> int:
    h = (C.c_uint8 * 64).from_buffer_copy(handle)
    p = _P()
    _check(lib().rt_ipc_open(device, h, C.byref(p)))
    return p.value


def ipc_close(device: int, ptr: int) -> None:
    _check(lib().rt_ipc_close(device, ptr))


def orbit_camera(k: int, n: int, width: int, height: int, radius: float = 60.0, cam_height: float = 9.0,
                 fov_deg: float = 40.0) -> SceneCamera:
    out = SceneCamera()
    _check(lib().rt_orbit_camera(k, n, radius, cam_height, fov_deg, width, height, C.byref(out)))
    return out


class Key(enum.IntFlag):
    """Keys of camera_movement_system (src/camera.rs:125-253); values are the RT_KEY_* masks of rt_abi.h."""
    W = 0x001
    S = 0x002
    A = 0x004
    D = 0x008
    ARROW_UP = 0x010
    ARROW_DOWN = 0x020
    ARROW_LEFT = 0x040
    ARROW_RIGHT = 0x080
    DIGIT1 = 0x100
    DIGIT2 = 0x200


def camera_movement_system(settings: CameraSettings, keys: int, dt: float) -> CameraSettings:
    """One frame of the reference's keyboard controller, in place (camera.rs:125-253)."""
    _check(lib().rt_camera_step(C.byref(settings), int(keys), dt))
    return settings


def camera_script(settings: CameraSettings, keys, dt: float, width: int, height: int, seeds=None):
    """Scripted fly-through: one controller frame + uniform derivation per entry of `keys`; returns the SceneCameras."""
    n = len(keys)
    key_arr = (C.c_uint32 * n)(*[int(k) for k in keys])
    seed_arr = (C.c_float * n)(*seeds) if seeds is not None else None
    out = (SceneCamera * n)()
    _check(lib().rt_camera_script(C.byref(settings), key_arr, n, dt, seed_arr, width, height, out))
    return [out[i].copy() for i in range(n)]


# ------------------------------------------------------------------------------------------------
# renderer
# ------------------------------------------------------------------------------------------------
class Renderer:
    """One device, one accumulation image (or a row window of it)."""

    def __init__(self, width: int, height: int, device: int = 0, row_begin: int = 0, row_count: int = None, bands=None):
        """bands = (band_rows, n_shards, index): an interleaved shard (rt_create_bands) instead of a contiguous window."""
        self._h = C.c_void_p()
        self.bands = bands
        if bands is not None:
            band_rows, n_shards, index = bands
            row_begin, row_count = index * band_rows, shard_bands(height, band_rows, n_shards, index)
            _check(lib().rt_create_bands(device, width, height, band_rows, n_shards, index, C.byref(self._h)))
        else:
            if row_count is None:
                row_count = height - row_begin
            _check(lib().rt_create_window(device, width, height, row_begin, row_count, C.byref(self._h)))
        self.width, self.height, self.row_begin, self.row_count, self.device = width, height, row_begin, row_count, device

    def image_rows(self) -> np.ndarray:
        """Image row of every window row, in window order (what read_accum / read_rgba8 return)."""
        r = np.arange(self.row_count)
        if self.bands is None:
            return self.row_begin + r
        band_rows, n_shards, index = self.bands
        return (r // band_rows * n_shards + index) * band_rows + r % band_rows

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().rt_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_scene(self, scene: SphereCollection):
        _check(lib().rt_set_scene(self._h, _ptr(scene.spheres), len(scene.spheres), scene.count))

    def init(self):
        _check(lib().rt_init(self._h))

    def update(self, camera: SceneCamera):
        _check(lib().rt_update(self._h, C.byref(camera)))

    def update_n(self, cameras):
        arr = (SceneCamera * len(cameras))(*cameras)
        _check(lib().rt_update_n(self._h, arr, len(cameras)))

    def synchronize(self):
        _check(lib().rt_synchronize(self._h))

    def read_accum(self, out: np.ndarray = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.row_count, self.width, 4), dtype=np.float32)
        _check(lib().rt_read_accum(self._h, _ptr(out)))
        return out

    def write_accum(self, accum: np.ndarray):
        accum = np.ascontiguousarray(accum, dtype=np.float32)
        assert accum.size == self.row_count * self.width * 4
        _check(lib().rt_write_accum(self._h, _ptr(accum)))

    def read_rgba8(self, out: np.ndarray = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.row_count, self.width, 4), dtype=np.uint8)
        _check(lib().rt_read_rgba8(self._h, _ptr(out)))
        return out

    def read_rgba8_into(self, host_ptr: int):
        _check(lib().rt_read_rgba8(self._h, C.c_void_p(host_ptr)))

    def read_rgba8_async_into(self, host_ptr: int):
        """Non-blocking read-back into a (pinned) host buffer; pair with read_wait()."""
        _check(lib().rt_read_rgba8_async(self._h, C.c_void_p(host_ptr)))

    def read_wait(self):
        _check(lib().rt_read_wait(self._h))

    def quantize(self):
        _check(lib().rt_quantize(self._h))

    def quantize_into(self, device_ptr: int):
        _check(lib().rt_quantize_into(self._h, C.c_void_p(device_ptr)))

    def read_accum_into(self, host_ptr: int):
        _check(lib().rt_read_accum(self._h, C.c_void_p(host_ptr)))

    def set_scene_from(self, host_ptr: int, buffer_len: int, count: int):
        _check(lib().rt_set_scene(self._h, C.c_void_p(host_ptr), buffer_len, count))

    def rays_traced(self) -> int:
        v = C.c_uint64(0)
        _check(lib().rt_rays_traced(self._h, C.byref(v)))
        return v.value

    def exact_tests(self) -> int:
        v = C.c_uint64(0)
        _check(lib().rt_exact_tests(self._h, C.byref(v)))
        return v.value

    def wavefront_counts(self):
        """Rays traced per bounce (index = bounce, >= 1) by the last deep update."""
        buf = (C.c_uint32 * 80)()
        n = C.c_uint32(0)
        _check(lib().rt_wavefront_counts(self._h, buf, 80, C.byref(n)))
        return [int(buf[i]) for i in range(n.value)]

    def set_rgba8_output(self, enable: bool = True):
        """Depth-1 updates also write the RGBA8 frame the next read_rgba8_async* call sends (no quantise kernel)."""
        _check(lib().rt_set_rgba8_output(self._h, 1 if enable else 0))

    def set_rgba8_target(self, device_ptr):
        """Fused gather: depth-1 updates store the window's RGBA8 texels at device_ptr (this GPU's memory or a peer's);
        None switches it off."""
        _check(lib().rt_set_rgba8_target(self._h, C.c_void_p(device_ptr or 0)))

    def stream_signal(self, device_flag: int, value: int):
        _check(lib().rt_stream_signal(self._h, C.c_void_p(device_flag), value))

    def stream_wait(self, device_flags: int, n: int, value: int, timeout_ms: int = 2000):
        _check(lib().rt_stream_wait(self._h, C.c_void_p(device_flags), n, value, timeout_ms))

    def send_rgba8_async(self, device_full_rgba8: int):
        """Pipelined gather: the window's RGBA8 rows into their places in a full image (this GPU's memory or a peer's) by a
        copy engine on the copy stream; pair with copy_signal / copy_wait and read_wait()."""
        _check(lib().rt_send_rgba8_async(self._h, C.c_void_p(device_full_rgba8)))

    def copy_signal(self, device_flag: int, value: int):
        _check(lib().rt_copy_signal(self._h, C.c_void_p(device_flag), value))

    def copy_wait(self, device_flags: int, n: int, value: int, timeout_ms: int = 2000):
        _check(lib().rt_copy_wait(self._h, C.c_void_p(device_flags), n, value, timeout_ms))

    def stream_wait_timed_out(self) -> int:
        v = C.c_uint32(0)
        _check(lib().rt_stream_wait_timed_out(self._h, C.byref(v)))
        return v.value

    def set_precision(self, precision: int):
        """RT_PRECISION_EXACT (default, bit-identical to the oracle), RT_PRECISION_FAST_SHADING (MUFU + FMA in scatter / sky /
        running mean, primary rays exact) or RT_PRECISION_FAST_ALL (ray generation too)."""
        _check(lib().rt_set_precision(self._h, precision))

    def tree_build_ms(self) -> float:
        v = C.c_double(0)
        _check(lib().rt_tree_build_ms(self._h, C.byref(v)))
        return v.value

    def tree_info(self):
        """(nodes, levels, bounce path) of the scene's tree: path 1 = shared-memory image, 2 = walked in global memory,
        3 = plain records streamed, 0 = no tree yet."""
        n, d, b = C.c_uint32(0), C.c_uint32(0), C.c_uint32(0)
        _check(lib().rt_tree_info(self._h, C.byref(n), C.byref(d), C.byref(b)))
        return n.value, d.value, b.value

    def reset_counters(self):
        _check(lib().rt_reset_counters(self._h))

    def timer_begin(self):
        _check(lib().rt_timer_begin(self._h))

    def timer_end(self) -> float:
        ms = C.c_float(0)
        _check(lib().rt_timer_end(self._h, C.byref(ms)))
        return ms.value

    def kernel_launches(self) -> int:
        v = C.c_uint64(0)
        _check(lib().rt_kernel_launches(self._h, C.byref(v)))
        return v.value

    def set_kernel_variant(self, variant: int):
        _check(lib().rt_set_kernel_variant(self._h, variant))

    def accum_device_ptr(self) -> int:
        p = C.c_void_p()
        _check(lib().rt_accum_device_ptr(self._h, C.byref(p)))
        return p.value

    def rgba8_device_ptr(self) -> int:
        p = C.c_void_p()
        _check(lib().rt_rgba8_device_ptr(self._h, C.byref(p)))
        return p.value

    def stream(self) -> int:
        p = C.c_void_p()
        _check(lib().rt_stream(self._h, C.byref(p)))
        return p.value or 0


def shard_bands(height: int, band_rows: int, n_shards: int, index: int) -> int:
    """Rows an interleaved shard owns — rt_shard_bands."""
    c = C.c_uint32(0)
    _check(lib().rt_shard_bands(height, band_rows, n_shards, index, C.byref(c)))
    return c.value


def shard_rows(height: int, n_shards: int, index: int):
    """(row_begin, row_count) of shard `index` — rt_shard_rows."""
    b, c = C.c_uint32(0), C.c_uint32(0)
    _check(lib().rt_shard_rows(height, n_shards, index, C.byref(b), C.byref(c)))
    return b.value, c.value


class MultiRenderer:
    """One image sharded by row blocks over several devices of this host (rt_multi_*)."""

    def __init__(self, devices, width: int, height: int, band_rows: int = 0):
        """band_rows > 0: interleaved bands of that many rows (balanced), else contiguous row blocks."""
        self._h = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        if band_rows:
            _check(lib().rt_multi_create_bands(arr, len(devices), width, height, band_rows, C.byref(self._h)))
        else:
            _check(lib().rt_multi_create(arr, len(devices), width, height, C.byref(self._h)))
        self.width, self.height, self.devices = width, height, list(devices)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().rt_multi_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_scene(self, scene: SphereCollection):
        _check(lib().rt_multi_set_scene(self._h, _ptr(scene.spheres), len(scene.spheres), scene.count))

    def init(self):
        _check(lib().rt_multi_init(self._h))

    def update(self, camera: SceneCamera):
        _check(lib().rt_multi_update(self._h, C.byref(camera)))

    def gather_rgba8(self, to_host: bool = True):
        out = np.empty((self.height, self.width, 4), dtype=np.uint8) if to_host else None
        _check(lib().rt_multi_gather_rgba8(self._h, _ptr(out) if to_host else None))
        return out

    def gather_accum(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        _check(lib().rt_multi_gather_accum(self._h, _ptr(out)))
        return out

    def rays_traced(self) -> int:
        v = C.c_uint64(0)
        _check(lib().rt_multi_rays_traced(self._h, C.byref(v)))
        return v.value


# ------------------------------------------------------------------------------------------------
# the render-graph node (src/lib.rs:326-422)
# ------------------------------------------------------------------------------------------------
class ComputeShaderState(enum.Enum):
    Loading = 0
    Init = 1
    Update = 2


class ComputeShaderNode:
    """Mirror of `ComputeShaderNode` (src/lib.rs:332-422): `update()` advances the state machine
    Loading -> Init -> Update(1) <-> Update(0) (lib.rs:345-377) and `run()` issues the dispatch of
    the current state (lib.rs:379-421): nothing while Loading, `init` in Init, `update` in Update.
    Pipelines are compiled ahead of time here, so Loading lasts exactly one `update()` call."""

    def __init__(self, renderer: Renderer):
        self.renderer = renderer
        self.state = ComputeShaderState.Loading
        self.index = 0   # the ping-pong index the reference carries; kept for API parity only

    def update(self):
        if self.state is ComputeShaderState.Loading:
            self.state = ComputeShaderState.Init
        elif self.state is ComputeShaderState.Init:
            self.state, self.index = ComputeShaderState.Update, 1
        else:
            self.index = 1 - self.index

    def run(self, camera: SceneCamera):
        if self.state is ComputeShaderState.Init:
            self.renderer.init()
        elif self.state is ComputeShaderState.Update:
            self.renderer.update(camera)


RT_STATIC_MAX_SPHERES = 4096


def update_static(device: int, device_spheres: int, n_spheres: int, camera: SceneCamera, width: int, height: int,
                  device_accum: int, stream: int = 0) -> None:
    """rt_update_static: one `update` over a whole image with DEVICE pointers for spheres and accumulation image, no
    renderer handle and no allocation (scratch in a static device array; at most RT_STATIC_MAX_SPHERES spheres; calls that
    overlap on one GPU race on that array).  Asynchronous on `stream` (0 = default stream)."""
    _check(lib().rt_update_static(device, C.c_void_p(device_spheres), n_spheres, C.byref(camera), width, height,
                                  C.c_void_p(device_accum), C.c_void_p(stream)))


def render(scene: SphereCollection, camera: SceneCamera, width: int, height: int, samples: int = 1,
           device: int = 0) -> np.ndarray:
    """Headless render(scene, camera, w, h): rgba32f (h, w, 4) = (mean rgb, sample count)."""
    out = np.empty((height, width, 4), dtype=np.float32)
    _check(lib().rt_render(device, _ptr(scene.spheres), scene.count, C.byref(camera), width, height, samples, _ptr(out)))
    return out


def render_frames(scene: SphereCollection, cameras, width: int, height: int, device: int = 0) -> np.ndarray:
    out = np.empty((len(cameras), height, width, 4), dtype=np.uint8)
    arr = (SceneCamera * len(cameras))(*cameras)
    _check(lib().rt_render_frames(device, _ptr(scene.spheres), scene.count, arr, len(cameras), width, height, _ptr(out)))
    return out


def write_png(path: str, rgba8: np.ndarray, gamma: float = None, alpha: bool = False) -> None:
    """Write an RGBA8 image (h, w, 4) as a PNG (8-bit RGB, or RGBA with alpha=True) using only zlib: the display path
    of the reference shows the Rgba32Float texture through a sprite (src/lib.rs:63-102); headless, the frame goes to a
    file instead.  gamma as in write_ppm."""
    import struct
    import zlib
    img = np.asarray(rgba8, dtype=np.uint8)
    img = img[..., :4] if alpha else img[..., :3]
    if gamma:
        rgb = np.round(255.0 * (img[..., :3].astype(np.float32) / 255.0) ** (1.0 / gamma)).astype(np.uint8)
        img = np.concatenate([rgb, img[..., 3:]], axis=-1) if alpha else rgb
    h, w, c = img.shape
    raw = np.empty((h, 1 + w * c), dtype=np.uint8)
    raw[:, 0] = 0                                   # filter type 0 (none) per scanline
    raw[:, 1:] = img.reshape(h, w * c)

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6 if alpha else 2, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)))
        f.write(chunk(b"IEND", b""))


def read_png(path: str) -> np.ndarray:
    """Inverse of write_png for the files it writes (8-bit RGB / RGBA, filter 0): (h, w, 3 or 4) uint8."""
    import struct
    import zlib
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h, c = 8, b"", 0, 0, 3
    while pos < len(data):
        n, tag = struct.unpack(">I", data[pos:pos + 4])[0], data[pos + 4:pos + 8]
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert depth == 8 and ctype in (2, 6)
            c = 4 if ctype == 6 else 3
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + w * c)
    assert not raw[:, 0].any(), "only filter type 0 is supported"
    return raw[:, 1:].reshape(h, w, c).copy()


def diff_image(a: np.ndarray, b: np.ndarray, scale: int = 32):
    """Parity debugging: per-pixel difference of two rgba32f accumulations (or RGBA8 frames).  Returns
    (rgba8 picture, report): the picture shows |a - b| of the RGBA8 quantisations times `scale` (clamped), pixels over
    the 1/255 tolerance in pure red; the report counts them (the mismatch-pixel count the port target asks for)."""
    def q(x):
        x = np.asarray(x)
        if x.dtype == np.uint8:
            return x[..., :3].astype(np.int16)
        return np.floor(np.clip(np.nan_to_num(x[..., :3].astype(np.float32), nan=0.0), 0.0, 1.0) * 255.0 + 0.5).astype(np.int16)
    d = np.abs(q(a) - q(b))
    worst = d.max(axis=-1)
    pic = np.zeros(d.shape[:2] + (4,), dtype=np.uint8)
    pic[..., :3] = np.clip(d * scale, 0, 255).astype(np.uint8)
    pic[worst > 1] = (255, 0, 0, 0)
    pic[..., 3] = 255
    report = {"pixels": int(worst.size), "pixels_differing": int((worst > 0).sum()), "pixels_over_1_255": int((worst > 1).sum()),
              "max_levels": int(worst.max()) if worst.size else 0}
    return pic, report


def write_ppm(path: str, rgba8: np.ndarray, gamma: float = None) -> None:
    """Write an RGBA8 image (h, w, 4) — e.g. Renderer.read_rgba8() — as a binary PPM for inspection.  The shader stores
    linear colour (the reference leaves any transfer curve to the display path, src/lib.rs:63-102); pass gamma=2.2 to
    apply one for viewing.  File output only: no windowing on this box."""
    img = np.asarray(rgba8)[..., :3]
    if gamma:
        img = np.round(255.0 * (img.astype(np.float32) / 255.0) ** (1.0 / gamma)).astype(np.uint8)
    h, w = img.shape[:2]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(img, dtype=np.uint8).tobytes())
