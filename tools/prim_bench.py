"""Developer timing of the primary pass (variant 4) on the bench scenes; RT_B200_LIB selects the build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402
from dev_bench import time_variant  # noqa: E402

if __name__ == "__main__":
    tag = os.environ.get("RT_B200_LIB", "default")
    W, H = 3840, 2160
    out = []
    for name, scene, cam, w, h, it in (
        ("r1024 d1", scenes.random_scene(1024), scenes.camera(W, H, 1, 1), W, H, 40),
        ("r1024 pinhole", scenes.random_scene(1024), scenes.camera(W, H, 1, 1, defocus_angle=0.0), W, H, 40),
        ("default d1", scenes.default_scene(), scenes.camera(W, H, 1, 1), W, H, 40),
        ("grid65536 1080p", scenes.grid_scene(65536), rt.orbit_camera(3, 256, 1920, 1080), 1920, 1080, 20),
        ("r1024 d30", scenes.random_scene(1024), scenes.camera(W, H, 30, 500), W, H, 10),
    ):
        ms, rays, exact = time_variant(scene, cam, w, h, 4, it)
        out.append(f"{name} {ms:.4f}")
    if "--deep" in sys.argv:   # 65,536 spheres with bounces: tree walked in global memory (4) against streamed records (5)
        cam = rt.orbit_camera(3, 256, 1920, 1080)
        cam.max_depth = 8.0
        for variant, it in ((4, 5), (5, 1)):
            ms, rays, exact = time_variant(scenes.grid_scene(65536), cam, 1920, 1080, variant, it)
            out.append(f"grid65536 d8 v{variant} {ms:.3f} ({rays / 1e6:.2f} Mrays)")
    print(os.path.basename(tag), " | ".join(out), flush=True)
