"""CPU tests of the oracle against the committed fixtures (tests/golden) and hand cases.
The reference has no tests for this path: the pins are the known-answer
values derived from the shader text by tests/golden/make_golden.py (pure Python, independent of
the oracle) and by the survey."""
import ctypes as C
import json
import math
import os

import numpy as np
import pytest

import gpu_ray_tracing_b200 as rt
import oracle_lib as ol
import scenes

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KAT = json.load(open(os.path.join(GOLDEN, "kat.json")))


def f32_bits(x):
    return int(np.float32(x).view(np.uint32))


def test_hash_known_answers():
    o = ol.oracle()
    for table in ("hash", "survey_hash"):
        for v, expect in KAT[table].items():
            assert o.oracle_hash(int(v)) == expect, (table, v)


def test_random_float_known_answers_and_range():
    o = ol.oracle()
    for table in ("random_float_bits", "survey_random_float_bits"):
        for v, bits in KAT[table].items():
            assert f32_bits(o.oracle_random_float(int(v))) == bits, (table, v)
    vals = np.array([o.oracle_random_float(v) for v in range(0, 200000, 37)], dtype=np.float32)
    assert vals.min() >= 0.0 and vals.max() <= 1.0
    assert abs(float(vals.mean()) - 0.5) < 0.02


def test_pinned_sincos_and_pow_against_libm():
    """include/rt_pinned_math.h sin, cos within 2 ulp (or 2^-22 absolute), pow5 within 3 ulp of the exact value over the shader ranges."""
    o = ol.oracle()
    s, c = C.c_float(), C.c_float()
    worst = 0.0
    for a in np.linspace(0.0, 6.2831855, 20001, dtype=np.float32).tolist() + [6.2831855, 3.1415927, 1.5707964, 4.712389]:
        o.oracle_sincos(C.c_float(a), C.byref(s), C.byref(c))
        af = float(np.float32(a))
        for got, want in ((s.value, math.sin(af)), (c.value, math.cos(af))):
            err = abs(got - want)
            ulp = float(np.spacing(np.float32(abs(want)))) if want != 0 else 1e-45
            assert err <= max(2.0 * ulp, 2.0 ** -22), (a, got, want)
            worst = max(worst, err)
        assert abs(s.value * s.value + c.value * c.value - 1.0) < 1e-6
    for x in np.linspace(0.0, 2.0, 4001, dtype=np.float32).tolist():
        got, want = o.oracle_pow5(C.c_float(x)), float(np.float32(x)) ** 5
        assert abs(got - want) <= 3.0 * float(np.spacing(np.float32(want))) + 1e-45
    assert ol.oracle(libm=True).oracle_uses_libm() == 1 and o.oracle_uses_libm() == 0


def _hit(sphere, origin, direction, tmin=0.001, tmax=3.4e35):
    sp = np.zeros(1, dtype=rt.SPHERE_DTYPE)
    sp[0]["position"], sp[0]["radius"] = sphere[:3], sphere[3]
    o, d = np.array(origin, np.float32), np.array(direction, np.float32)
    out = np.zeros(8, dtype=np.float32)
    hit = ol.oracle().oracle_sphere_hit(ol._p(sp), ol._p(o), ol._p(d), C.c_float(tmin), C.c_float(tmax), ol._p(out))
    return bool(hit), out


def test_sphere_hit_hand_cases():
    # straight on: unit sphere at z=-5, ray down -z: roots 4 and 6
    hit, r = _hit((0, 0, -5, 1), (0, 0, 0), (0, 0, -1))
    assert hit and r[0] == 4.0 and tuple(r[1:4]) == (0, 0, -4) and tuple(r[4:7]) == (0, 0, 1) and r[7] == 1.0
    # miss: discriminant < 0
    assert not _hit((0, 0, -5, 1), (0, 0, 0), (0, 1, 0))[0]
    # inside the sphere: near root negative -> far root, back face, normal flipped toward the ray origin
    hit, r = _hit((0, 0, 0, 2), (0, 0, 0), (0, 0, -1))
    assert hit and r[0] == 2.0 and r[7] == 0.0 and tuple(r[4:7]) == (0, 0, 1)
    # un-normalised direction: t is in units of |d|
    hit, r = _hit((0, 0, -5, 1), (0, 0, 0), (0, 0, -2))
    assert hit and r[0] == 2.0
    # bounds are closed on rejection: root == tmin and root == tmax are both rejected (wgsl:196,198)
    assert not _hit((0, 0, -5, 1), (0, 0, 0), (0, 0, -1), tmin=6.0)[0]          # 4 <= 6, 6 <= 6
    hit, r = _hit((0, 0, -5, 1), (0, 0, 0), (0, 0, -1), tmin=4.0)
    assert hit and r[0] == 6.0                                                  # near root rejected, far accepted
    assert not _hit((0, 0, -5, 1), (0, 0, 0), (0, 0, -1), tmax=4.0)[0]          # 4 <= 4 rejected, 6 too far
    # sphere behind the ray
    assert not _hit((0, 0, 5, 1), (0, 0, 0), (0, 0, -1))[0]
    # tangent ray: discriminant exactly 0 counts as a hit
    hit, r = _hit((1, 0, -5, 1), (0, 0, 0), (0, 0, -1))
    assert hit and r[0] == 5.0


def test_sphere_list_tie_keeps_lowest_index_and_nearest_wins():
    sp = np.zeros(4, dtype=rt.SPHERE_DTYPE)
    sp["radius"] = 1.0
    sp[0]["position"], sp[1]["position"], sp[2]["position"], sp[3]["position"] = (0, 0, -9), (0, 0, -5), (0, 0, -5), (0, 0, -7)
    o, d = np.zeros(3, np.float32), np.array([0, 0, -1], np.float32)
    out, idx = np.zeros(8, np.float32), C.c_int32(-2)
    hit = ol.oracle().oracle_sphere_list_hit(ol._p(sp), 4, ol._p(o), ol._p(d), C.c_float(0.001), C.c_float(3.4e35), ol._p(out), C.byref(idx))
    assert hit and out[0] == 4.0 and idx.value == 1      # 1 and 2 tie at t=4: strict rejection keeps index 1
    hit = ol.oracle().oracle_sphere_list_hit(ol._p(sp), 0, ol._p(o), ol._p(d), C.c_float(0.001), C.c_float(3.4e35), ol._p(out), C.byref(idx))
    assert not hit and idx.value == -1


def test_camera_derivation_matches_independent_numpy_and_survey():
    for (w, h) in ((1280, 720), (3840, 2160)):
        ref = KAT[f"default_camera_{w}x{h}"]
        s = rt.CameraSettings.default()
        for cam in (rt.SceneCamera.from_settings(s, w, h, 0.5),):
            ocam = rt.SceneCamera()
            ol.oracle().oracle_camera_from_settings(C.byref(s), w, h, C.c_float(0.5), C.byref(ocam))
            assert cam.to_bytes() == ocam.to_bytes()          # product host helper == oracle restatement, bit for bit
            def close(a, b):
                return np.allclose(np.array(a, np.float32), np.array(b, np.float32), rtol=3e-7, atol=1e-9)
            assert close(cam.viewport_height, ref["viewport_height"]) and close(cam.viewport_width, ref["viewport_width"])
            assert close(cam.pixel_delta_u.tuple(), ref["pixel_delta_u"]) and close(cam.pixel_delta_v.tuple(), ref["pixel_delta_v"])
            assert close(cam.viewport_upper_left.tuple(), ref["viewport_upper_left"])
            assert close(cam.defocus_radius, ref["defocus_radius"])
            assert close(cam.defocus_disk_u.tuple(), ref["defocus_disk_u"]) and close(cam.defocus_disk_v.tuple(), ref["defocus_disk_v"])
    # survey KAT, 1280x720
    cam = rt.SceneCamera.from_settings(rt.CameraSettings.default(), 1280, 720, 0.0)
    assert abs(cam.viewport_height - 3.5265396) < 1e-6 and abs(cam.viewport_width - 6.269404) < 1e-6
    assert np.allclose(cam.viewport_upper_left.tuple(), (2.4041812, 2.2612867, 3.771899), atol=2e-6)
    assert abs(cam.defocus_radius - 0.052360356) < 1e-8
    assert cam.max_depth == 30.0 and cam.samples_per_pixel == 500.0 and cam.camera_has_moved == 1.0


def test_oracle_images_match_committed_fixtures():
    """The oracle reproduces tests/golden/images.npz bit for bit (fixtures made by make_golden.py)."""
    gold = np.load(os.path.join(GOLDEN, "images.npz"))
    cases = {
        "default_96x54_d30_f2": (scenes.default_scene(), 96, 54, 30, 500, (0.5, 0.25)),
        "default_64x36_d1": (scenes.default_scene(), 64, 36, 1, 1, (0.5,)),
        "random1024_80x45_d6": (scenes.random_scene(1024), 80, 45, 6, 500, (0.125,)),
        "grid4096_48x27_d3": (scenes.grid_scene(4096), 48, 27, 3, 500, (0.75,)),
    }
    for name, (scene, W, H, depth, spp, seeds) in cases.items():
        accum = np.zeros((H, W, 4), dtype=np.float32)
        for i, s in enumerate(seeds):
            cam = scenes.camera(W, H, max_depth=depth, spp=spp, random_seed=s, moved=1 if i == 0 else 0)
            accum, _ = ol.oracle_update(cam, scene, W, H, accum=accum)
        assert np.array_equal(accum.view(np.uint32), gold[name]), name


def test_update_semantics():
    """Reset, sample cap, running mean and the depth-0 / empty-scene sky (compute_shader.wgsl:339-363, 293-296)."""
    W, H = 16, 8
    scene = scenes.default_scene()
    start = np.full((H, W, 4), 0.25, dtype=np.float32)
    start[..., 3] = 3.0
    # cap reached and no reset: texel unchanged, no rays
    out, rays = ol.oracle_update(scenes.camera(W, H, 5, 3, moved=0), scene, W, H, accum=start.copy())
    assert rays == 0 and np.array_equal(out, start)
    # reset zeroes mean and count before sampling
    out, rays = ol.oracle_update(scenes.camera(W, H, 5, 3, moved=1), scene, W, H, accum=start.copy())
    assert (out[..., 3] == 1.0).all() and rays >= W * H
    # running mean: second sample averages
    c1 = scenes.camera(W, H, 1, 10, random_seed=0.5, moved=1)
    c2 = scenes.camera(W, H, 1, 10, random_seed=0.75, moved=0)
    a1, _ = ol.oracle_update(c1, scene, W, H)
    only2, _ = ol.oracle_update(scenes.camera(W, H, 1, 10, random_seed=0.75, moved=1), scene, W, H)
    a2, _ = ol.oracle_update(c2, scene, W, H, accum=a1.copy())
    assert (a2[..., 3] == 2.0).all()
    # the second frame's seed depends on samples_until_now, so compare against the definition instead
    assert not np.array_equal(a2[..., :3], a1[..., :3]) and only2.shape == a1.shape
    # max_depth 0: pure sky gradient, no traversal
    sky, rays = ol.oracle_update(scenes.camera(W, H, 0, 1), scene, W, H)
    assert rays == 0 and (sky[..., 2] == 1.0).all() and (sky[..., 0] >= 0.5).all() and (sky[..., 0] <= 1.0).all()
    # row windows are slices of the full image (absolute pixel coordinates in the seed)
    full, _ = ol.oracle_update(scenes.camera(W, H, 4, 2), scene, W, H)
    part, _ = ol.oracle_update(scenes.camera(W, H, 4, 2), scene, W, H, row_begin=3, row_count=4)
    assert np.array_equal(part, full[3:7])


def test_libm_oracle_stays_within_tolerance_of_pinned_oracle():
    """How far does an implementation with different sin/cos/pow rounding land?  (report + bound)"""
    W, H = 160, 90
    scene, cam = scenes.default_scene(), scenes.camera(W, H, 30, 500)
    a, _ = ol.oracle_update(cam, scene, W, H)
    b, _ = ol.oracle_update(cam, scene, W, H, libm=True)
    rep = ol.compare_images(a, b)
    print("pinned vs libm oracle:", rep)
    assert rep["pixels_over_1_255"] <= 0.02 * rep["pixels"]
    d1a, _ = ol.oracle_update(scenes.camera(W, H, 1, 1, defocus_angle=0.0), scene, W, H)
    d1b, _ = ol.oracle_update(scenes.camera(W, H, 1, 1, defocus_angle=0.0), scene, W, H, libm=True)
    assert np.array_equal(d1a, d1b)     # no transcendental on the depth-1 pinhole path
