"""Developer timing script (not the contract bench): variants side by side on one GPU."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402


def ffma_peak():
    lib = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libffma_peak.so"))
    tf, sms = C.c_double(0), C.c_int(0)
    rc = lib.ffma_peak(0, 10, C.byref(tf), None, C.byref(sms))
    return rc, tf.value, sms.value


def time_variant(scene, cam, W, H, variant, iters):
    with rt.Renderer(W, H) as r:
        r.set_kernel_variant(variant)
        r.set_scene(scene)
        r.init()
        for _ in range(3):
            r.update(cam)
        r.synchronize()
        r.reset_counters()
        r.timer_begin()
        for _ in range(iters):
            r.update(cam)
        ms = r.timer_end() / iters
        rays = r.rays_traced() / iters
        exact = r.exact_tests() / iters
    return ms, rays, exact


if __name__ == "__main__":
    print("ffma_peak", ffma_peak(), flush=True)
    scene5, cam5 = scenes.grid_scene(65536), rt.orbit_camera(3, 256, 1920, 1080)
    for variant in (0, 3):
        ms, rays, exact = time_variant(scene5, cam5, 1920, 1080, variant, 3)
        print(f"grid65536 1080p d1     variant {variant}: {ms:8.3f} ms  {rays / ms / 1e3:9.1f} Mrays/s  "
              f"{rays * 18 * 65536 / ms / 1e9:7.2f} TFLOP/s(alg)  exact/ray {exact / max(rays, 1):.2f}", flush=True)
    for variant in (4, 2):
        ms, rays, exact = time_variant(scenes.default_scene(), scenes.camera(1280, 720, 30, 500), 1280, 720, variant, 20)
        print(f"default 720p depth30   variant {variant}: {ms:8.3f} ms  {rays / ms / 1e3:9.1f} Mrays/s  rays/frame {rays:.0f}", flush=True)
    W, H = 3840, 2160
    for name, scene, cam in (
        ("random1024 depth1", scenes.random_scene(1024), scenes.camera(W, H, 1, 1)),
        ("random1024 d1 pinhole", scenes.random_scene(1024), scenes.camera(W, H, 1, 1, defocus_angle=0.0)),
        ("default depth1", scenes.default_scene(), scenes.camera(W, H, 1, 1)),
        ("random1024 depth30", scenes.random_scene(1024), scenes.camera(W, H, 30, 500)),
    ):
        for variant in (4, 2):
            iters = 20 if variant != 1 else 3
            ms, rays, exact = time_variant(scene, cam, W, H, variant, iters)
            n = scene.count
            print(f"{name:22s} variant {variant}: {ms:8.3f} ms  {rays / ms / 1e3:9.1f} Mrays/s  "
                  f"{rays * 18 * n / ms / 1e9:7.2f} TFLOP/s(alg)  exact/ray {exact / max(rays, 1):.2f}", flush=True)
    print("ffma_peak", ffma_peak(), flush=True)
