"""Regenerates the committed fixtures.  The reference holds no golden vectors for this path
(SURVEY.md 8c), so these are produced by this project:
  kat.json ............ integer-hash / random_float / layout / default-camera known answers, derived
                        from the shader text (compute_shader.wgsl:50-63) and camera.rs:293-351 by an
                        independent pure-Python / numpy evaluation in THIS script (not by the oracle)
  images.npz .......... small oracle renders (bit patterns) that pin the oracle against regressions
                        and give the GPU box fixtures that do not need /root/reference or a rebuild
Run: python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

M = 0xFFFFFFFF


def py_hash(s):
    """compute_shader.wgsl:50-59, shift binding tighter than xor (naga)."""
    s ^= 2747636419
    s = (s * 2654435769) & M
    s ^= s >> 16
    s = (s * 2654435769) & M
    s ^= s >> 16
    s = (s * 2654435769) & M
    return s


def py_random_float_bits(v):
    """f32(hash(v)) / f32(4294967295.0): the literal rounds to 2^32, the u32->f32 conversion rounds to nearest even."""
    f = np.float32(np.uint32(py_hash(v))) / np.float32(4294967296.0)
    return int(np.float32(f).view(np.uint32))


def default_camera_f32(width, height):
    """camera.rs:293-351 in numpy float32, glam semantics (normalize = v * (1/len))."""
    f = np.float32
    fov, focus, defocus = f(20.0), f(10.0), f(0.6)
    look_from, look_at, vup = np.array([13, 2, 3], f), np.array([0, 0, 0], f), np.array([0, 1, 0], f)
    rads = f(0.017453292519943295)
    aspect = f(width) / f(height)
    theta = fov * rads
    h = f(np.tan(np.float64(theta / f(2.0))))          # correctly rounded tan of the f32 argument
    vh = f(2.0) * h * focus
    vw = vh * aspect

    def dot(a, b):
        return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]

    def norm(a):
        return a * (f(1.0) / np.sqrt(dot(a, a)))

    def cross(a, b):
        return np.array([a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]], f)

    w = norm(look_from - look_at)
    u = norm(cross(vup, w))
    v = cross(w, u)
    vu, vv = vw * u, (-vh) * v
    pdu, pdv = vu / f(width), vv / f(height)
    vul = look_from - focus * w - vu / f(2.0) - vv / f(2.0)
    dr = focus * f(np.tan(np.float64((defocus / f(2.0)) * rads)))
    return {"viewport_height": float(vh), "viewport_width": float(vw), "w": w.tolist(), "u": u.tolist(), "v": v.tolist(),
            "pixel_delta_u": pdu.tolist(), "pixel_delta_v": pdv.tolist(), "viewport_upper_left": vul.tolist(),
            "defocus_radius": float(dr), "defocus_disk_u": (u * dr).tolist(), "defocus_disk_v": (v * dr).tolist()}


def main():
    kat = {
        "hash": {str(v): py_hash(v) for v in [0, 1, 2, 3, 73, 51, 1000, 0xFFFFFFFF, 0xDEADBEEF, 25, 0x80000000, 12345678]},
        "survey_hash": {"0": 0x67b2772f, "1": 0x08fcaab9, "2": 0x55626429, "3": 0xeb7ef346, "73": 0x569dc710,
                        "51": 0xd908a1bc, "1000": 0x25089dd7, "4294967295": 0xc846935b, "3735928559": 0xd4e651b1},
        "random_float_bits": {str(v): py_random_float_bits(v) for v in [0, 1, 3, 7, 1000, 0xFFFFFFFF, 424242]},
        "survey_random_float_bits": {"0": 0x3ecf64ee, "1": 0x3d0fcaac, "3": 0x3f6b7ef3},
        "layout": {"SceneCamera": 176, "GpuSphere": 32, "CameraSettings": 60,
                   "camera_offsets": {"center": 0, "viewport_height": 12, "viewport_upper_left": 16, "viewport_width": 28,
                                      "pixel_delta_u": 32, "defocus_angle": 44, "pixel_delta_v": 48, "aspect_ratio": 60,
                                      "defocus_disk_u": 64, "_padding0": 76, "viewport_u": 80, "_padding1": 92,
                                      "defocus_disk_v": 96, "max_depth": 108, "look_from": 112, "samples_per_pixel": 124,
                                      "look_at": 128, "camera_has_moved": 140, "vup": 144, "random_seed": 156,
                                      "viewport_v": 160, "defocus_radius": 172}},
        "default_camera_1280x720": default_camera_f32(1280, 720),
        "default_camera_3840x2160": default_camera_f32(3840, 2160),
    }
    with open(os.path.join(HERE, "kat.json"), "w") as fh:
        json.dump(kat, fh, indent=1)

    import oracle_lib as ol
    import scenes
    images = {}
    for name, scene, W, H, depth, spp, seeds in (
        ("default_96x54_d30_f2", scenes.default_scene(), 96, 54, 30, 500, (0.5, 0.25)),
        ("default_64x36_d1", scenes.default_scene(), 64, 36, 1, 1, (0.5,)),
        ("random1024_80x45_d6", scenes.random_scene(1024), 80, 45, 6, 500, (0.125,)),
        ("grid4096_48x27_d3", scenes.grid_scene(4096), 48, 27, 3, 500, (0.75,)),
    ):
        accum = np.zeros((H, W, 4), dtype=np.float32)
        for i, s in enumerate(seeds):
            cam = scenes.camera(W, H, max_depth=depth, spp=spp, random_seed=s, moved=1 if i == 0 else 0)
            accum, _ = ol.oracle_update(cam, scene, W, H, accum=accum)
        images[name] = accum.view(np.uint32).copy()
    np.savez_compressed(os.path.join(HERE, "images.npz"), **images)
    print("wrote kat.json and images.npz:", {k: v.shape for k, v in images.items()})


if __name__ == "__main__":
    main()
