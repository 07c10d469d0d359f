"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol
include/rt_abi.h declares, keeps the reference's byte layouts, fails loudly without a device and
the host-side helpers (scenes, camera, node state machine, row sharding) behave like the reference."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

import gpu_ray_tracing_b200 as rt
from gpu_ray_tracing_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KAT = json.load(open(os.path.join(ROOT, "tests", "golden", "kat.json")))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "rt_abi.h")).read()
    return sorted(set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 40
    handle = C.CDLL(rt.LIB_PATH)
    missing = [n for n in names if not hasattr(handle, n)]
    assert not missing, missing
    unbound = [n for n in names if n not in api._SIGNATURES]
    assert not unbound, f"declared in rt_abi.h but not bound in api.py: {unbound}"
    rt.lib()


def test_layouts_match_reference_structs():
    lay = KAT["layout"]
    assert C.sizeof(rt.SceneCamera) == lay["SceneCamera"] and C.sizeof(rt.GpuSphere) == lay["GpuSphere"]
    assert C.sizeof(rt.CameraSettings) == lay["CameraSettings"]
    for field, off in lay["camera_offsets"].items():
        assert getattr(rt.SceneCamera, field).offset == off, field
    assert rt.GpuSphere.radius.offset == 12 and rt.GpuSphere.color.offset == 16
    assert rt.SPHERE_DTYPE.fields["radius"][1] == 12 and rt.SPHERE_DTYPE.fields["color"][1] == 16
    cam = rt.SceneCamera.from_settings(rt.CameraSettings.default(), 1280, 720, 0.5)
    raw = np.frombuffer(cam.to_bytes(), dtype="<f4")
    assert raw.shape == (44,) and tuple(raw[0:3]) == (13.0, 2.0, 3.0) and raw[11] == np.float32(0.6)
    assert raw[27] == 30.0 and raw[31] == 500.0 and raw[35] == 1.0 and raw[39] == 0.5
    assert rt.SceneCamera.from_bytes(cam.to_bytes()).to_bytes() == cam.to_bytes()


@pytest.mark.skipif(rt.device_count() > 0, reason="checks the no-device error path")
def test_no_device_fails_loudly_not_silently():
    with pytest.raises(rt.RtError) as e:
        rt.Renderer(64, 64)
    assert e.value.code == api.RT_ERR_NO_DEVICE
    scene = rt.create_default_spheres(1)
    cam = rt.SceneCamera.from_settings(rt.CameraSettings.default(), 8, 8, 0.5)
    with pytest.raises(rt.RtError):
        rt.render(scene, cam, 8, 8)
    with pytest.raises(rt.RtError):
        rt.MultiRenderer([0, 1], 64, 64)


def test_argument_validation_codes():
    lib = rt.lib()
    assert lib.rt_update(None, None) == api.RT_ERR_INVALID_ARG
    assert b"null" in lib.rt_last_error()
    assert lib.rt_destroy(None) == api.RT_OK
    h = C.c_void_p()
    assert lib.rt_create_window(0, 0, 10, 0, 10, C.byref(h)) == api.RT_ERR_INVALID_ARG
    assert lib.rt_create_window(0, 10, 10, 8, 5, C.byref(h)) == api.RT_ERR_INVALID_ARG
    assert lib.rt_camera_from_settings(None, 1, 1, 0.0, None) == api.RT_ERR_INVALID_ARG
    b, c = C.c_uint32(), C.c_uint32()
    assert lib.rt_shard_rows(10, 0, 0, C.byref(b), C.byref(c)) == api.RT_ERR_INVALID_ARG
    assert lib.rt_shard_rows(3, 4, 0, C.byref(b), C.byref(c)) == api.RT_ERR_INVALID_ARG


def test_default_scene_recipe():
    """sphere.rs:45-153: ground + up to 196 small + 3 big; materials by color.w; pad floor 100."""
    for seed in (0, 7, 12345):
        sc = rt.create_default_spheres(seed)
        s = sc.spheres[: sc.count]
        assert 150 <= sc.count <= 200 and len(sc.spheres) >= max(sc.count, 100)
        assert tuple(s[0]["position"]) == (0, -1000, 0) and s[0]["radius"] == 1000 and tuple(s[0]["color"]) == (0.5, 0.5, 0.5, -2)
        assert tuple(s[-3]["position"]) == (0, 1, 0) and tuple(s[-3]["color"]) == (1.5, 0, 0, 2)
        assert tuple(s[-2]["position"]) == (-4, 1, 0) and np.allclose(s[-2]["color"], (0.4, 0.2, 0.1, -2))
        assert tuple(s[-1]["position"]) == (4, 1, 0) and np.allclose(s[-1]["color"], (0.7, 0.6, 0.5, 0))
        small = s[1:-3]
        assert (small["radius"] == np.float32(0.2)).all() and (small["position"][:, 1] == np.float32(0.2)).all()
        d = np.linalg.norm(small["position"] - np.array([4, 0.2, 0], np.float32), axis=1)
        assert (d > 0.9).all()
        w = small["color"][:, 3]
        diffuse, glass = w == -2.0, w == 2.0
        metal = ~diffuse & ~glass
        assert 0.65 < diffuse.mean() < 0.92 and (w[metal] >= 0).all() and (w[metal] < 0.5).all()
        assert (small["color"][metal][:, :3] >= 0.5).all() and (small["color"][glass][:, 0] == 1.5).all()
    assert rt.create_default_spheres(7).to_bytes() == rt.create_default_spheres(7).to_bytes()
    assert rt.create_default_spheres(7).to_bytes() != rt.create_default_spheres(8).to_bytes()


def test_benchmark_scenes():
    sc = rt.create_random_spheres(1024, 1234)
    assert sc.count == 1024 and sc.spheres[0]["radius"] == 1000
    r = sc.spheres["radius"][1:]
    assert (r >= 0.1).all() and (r <= 0.4).all() and (sc.spheres["position"][1:, 1] == r).all()
    assert (np.abs(sc.spheres["position"][1:, [0, 2]]) <= 11).all()
    g = rt.create_grid_spheres(65536, 4321)
    assert g.count == 65536 and (g.spheres["radius"][1:] == np.float32(0.2)).all()
    assert np.abs(g.spheres["position"][1:, 0]).max() <= 129
    cams = [rt.orbit_camera(k, 256, 1920, 1080) for k in (0, 64, 255)]
    assert cams[0].defocus_angle == 0.0 and cams[0].max_depth == 1.0 and cams[1].random_seed == 0.25
    assert abs(np.hypot(cams[1].center.x, cams[1].center.z) - 60.0) < 1e-3


def test_shard_rows_partition():
    for height in (8, 131, 1080, 2160, 4320):
        for n in (1, 2, 3, 4, 8):
            cover, prev_end = 0, 0
            for i in range(n):
                b, c = rt.shard_rows(height, n, i)
                assert b == prev_end and c > 0
                if i + 1 < n and (height + 3) // 4 >= n:
                    assert b % 4 == 0 and c % 4 == 0
                prev_end = b + c
                cover += c
            assert cover == height and prev_end == height


class FakeRenderer:
    def __init__(self):
        self.calls = []

    def init(self):
        self.calls.append("init")

    def update(self, camera):
        self.calls.append("update")


def test_compute_shader_node_state_machine():
    """lib.rs:345-421: Loading -> Init -> Update(1) <-> Update(0); run() dispatches per state."""
    r = FakeRenderer()
    node = rt.ComputeShaderNode(r)
    node.run(None)                       # Loading: nothing dispatched
    assert r.calls == [] and node.state is rt.ComputeShaderState.Loading
    states = []
    for _ in range(5):
        node.update()
        node.run(None)
        states.append((node.state.name, node.index))
    assert r.calls == ["init", "update", "update", "update", "update"]
    assert states == [("Init", 0), ("Update", 1), ("Update", 0), ("Update", 1), ("Update", 0)]


def test_write_ppm_roundtrip(tmp_path):
    img = (np.arange(5 * 7 * 4, dtype=np.uint32) % 251).astype(np.uint8).reshape(5, 7, 4)
    path = tmp_path / "x.ppm"
    rt.write_ppm(str(path), img)
    data = path.read_bytes()
    assert data.startswith(b"P6\n7 5\n255\n")
    assert data[len(b"P6\n7 5\n255\n"):] == img[..., :3].tobytes()


# ------------------------------------------------------------------------------------------------
# keyboard controller restatement (camera.rs:125-253), SURVEY.md section 8(f) row 3
# ------------------------------------------------------------------------------------------------
def _v(v):
    return np.array(v.tuple(), dtype=np.float64)


def test_camera_step_translations_follow_reference_axes():
    K = rt.Key
    dt = 0.25
    base = rt.CameraSettings.default()
    f0, a0 = _v(base.look_from), _v(base.look_at)
    forward = (f0 - a0) / np.linalg.norm(f0 - a0)          # from the target TO the camera (camera.rs:135-136)
    right = np.cross(forward, [0.0, 1.0, 0.0])
    right /= np.linalg.norm(right)
    cases = {K.W: forward * 2.0 * dt, K.S: -forward * 2.0 * dt, K.A: -right * 2.0 * dt, K.D: right * 2.0 * dt,
             K.ARROW_UP: np.array([0.0, dt, 0.0]), K.ARROW_DOWN: np.array([0.0, -dt, 0.0])}
    for key, delta in cases.items():
        s = rt.CameraSettings.default()
        s.camera_has_moved = 0
        rt.camera_movement_system(s, key, dt)
        np.testing.assert_allclose(_v(s.look_from), f0 + delta, rtol=0, atol=2e-6)
        np.testing.assert_array_equal(_v(s.look_at), a0)     # keys never move the target
        assert s.camera_has_moved == 1
    s = rt.CameraSettings.default()
    rt.camera_movement_system(s, K.W | K.S | K.A | K.D, dt)  # opposite keys cancel (up to rounding) but count as movement
    np.testing.assert_allclose(_v(s.look_from), f0, rtol=0, atol=4e-6)
    assert s.camera_has_moved == 1


def test_camera_step_flag_clears_on_first_idle_frame():
    s = rt.CameraSettings.default()
    assert s.camera_has_moved == 1                          # default starts with the reset flag on (camera.rs:35)
    before = _v(s.look_from)
    rt.camera_movement_system(s, 0, 0.016)
    assert s.camera_has_moved == 0
    np.testing.assert_array_equal(_v(s.look_from), before)
    rt.camera_movement_system(s, rt.Key.ARROW_LEFT, 0.016)
    assert s.camera_has_moved == 1
    rt.camera_movement_system(s, 0, 0.016)
    assert s.camera_has_moved == 0


def test_camera_step_yaw_is_applied_twice_per_frame_and_keeps_distance():
    # the arrow block is duplicated in the reference (camera.rs:166-186 and :187-207): one frame yaws by 2 * dt
    s = rt.CameraSettings.default()
    f0, a0 = _v(s.look_from), _v(s.look_at)
    dt = 0.1
    rt.camera_movement_system(s, rt.Key.ARROW_LEFT, dt)
    ang = 2.0 * dt
    c, sn = np.cos(ang), np.sin(ang)
    v = f0 - a0
    want = a0 + np.array([c * v[0] + sn * v[2], v[1], -sn * v[0] + c * v[2]])   # rotation about +Y by +ang
    np.testing.assert_allclose(_v(s.look_from), want, rtol=0, atol=5e-6)
    rt.camera_movement_system(s, rt.Key.ARROW_RIGHT, dt)
    np.testing.assert_allclose(_v(s.look_from), f0, rtol=0, atol=1e-5)
    # a full orbit: 2 pi / (2 dt) frames come back to the start, radius and height kept all the way
    n = 200
    dt = float(np.float32(2.0 * np.pi / (2.0 * n)))
    r0 = np.linalg.norm(f0 - a0)
    for _ in range(n):
        rt.camera_movement_system(s, rt.Key.ARROW_LEFT, dt)
        assert abs(np.linalg.norm(_v(s.look_from) - a0) - r0) < 2e-4
        assert abs(s.look_from.y - f0[1]) < 2e-5
    np.testing.assert_allclose(_v(s.look_from), f0, rtol=0, atol=2e-3)


def test_camera_step_pitch_stops_before_the_pole():
    s = rt.CameraSettings.default()
    a0 = _v(s.look_at)
    r0 = np.linalg.norm(_v(s.look_from) - a0)
    y_prev = s.look_from.y
    rose = 0
    for _ in range(400):                                    # far more than needed to reach the pole
        rt.camera_movement_system(s, rt.Key.DIGIT1, 0.02)
        d = (_v(s.look_from) - a0) / np.linalg.norm(_v(s.look_from) - a0)
        assert abs(d[1]) < 0.95                             # camera.rs:221, 238
        assert abs(np.linalg.norm(_v(s.look_from) - a0) - r0) < 5e-4
        rose += s.look_from.y != y_prev
        y_prev = s.look_from.y
    assert 10 < rose < 400                                  # it moved for a while, then every frame was refused
    assert s.camera_has_moved == 0                          # a refused pitch is not movement
    d = (_v(s.look_from) - a0) / r0
    assert abs(d[1]) > 0.9


def test_camera_script_matches_stepping_by_hand():
    W, H, n, dt = 64, 48, 12, 1.0 / 60.0
    keys = [rt.Key.ARROW_LEFT if f % 3 else rt.Key.W | rt.Key.ARROW_UP for f in range(n)]
    keys[5] = 0
    seeds = [f / 16.0 for f in range(n)]
    cams = rt.camera_script(rt.CameraSettings.default(), keys, dt, W, H, seeds=seeds)
    s = rt.CameraSettings.default()
    for f in range(n):
        rt.camera_movement_system(s, keys[f], dt)
        want = rt.SceneCamera.from_settings(s, W, H, random_seed=seeds[f])
        assert cams[f].to_bytes() == want.to_bytes()
        assert cams[f].camera_has_moved == (0.0 if f == 5 else 1.0)


def test_png_round_trip_and_diff_image(tmp_path):
    """PNG writer (zlib only), its reader, and the diff-image report."""
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (37, 53, 4), dtype=np.uint8)
    p = str(tmp_path / "x.png")
    rt.write_png(p, img)
    assert np.array_equal(rt.read_png(p), img[..., :3])
    rt.write_png(p, img, alpha=True)
    assert np.array_equal(rt.read_png(p), img)
    a = rng.random((20, 30, 4), dtype=np.float32)
    b = a.copy()
    b[3, 4, 0] += 3.0 / 255.0          # over the tolerance
    b[5, 6, 1] += 1.0 / 255.0          # one level: within it
    pic, rep = rt.diff_image(a, b)
    assert rep["pixels"] == 600 and rep["pixels_over_1_255"] == 1 and rep["pixels_differing"] == 2
    assert tuple(pic[3, 4]) == (255, 0, 0, 255) and pic[0, 0, :3].sum() == 0


def test_committed_scene_bytes_match_the_generator():
    """tests/golden/random1024_seed1234.bin (what the CPU arms of bench.py read) == rt_create_random_spheres(1234, 1024)."""
    raw = open(os.path.join(ROOT, "tests", "golden", "random1024_seed1234.bin"), "rb").read()
    assert raw == rt.create_random_spheres(1024, 1234).to_bytes()


def test_band_shards_partition_the_rows():
    """rt_shard_bands: interleaved shards own every row exactly once, for ragged heights and more shards than bands."""
    for H, band, n in ((4320, 16, 8), (203, 8, 3), (64, 8, 8), (17, 8, 2), (100, 24, 5)):
        total = sum(rt.shard_bands(H, band, n, i) for i in range(n))
        assert total == H, (H, band, n, total)
    assert rt.shard_bands(16, 8, 4, 3) == 0
    c = C.c_uint32()
    assert rt.lib().rt_shard_bands(100, 12, 2, 0, C.byref(c)) == api.RT_ERR_INVALID_ARG      # not a multiple of 8
    assert rt.lib().rt_shard_bands(100, 8, 2, 2, C.byref(c)) == api.RT_ERR_INVALID_ARG
