"""Fixed workload for ncu: the config-5 frame (65,536 grid spheres, 1920x1080, orbit pose 3 of 256), depth argv[1]."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

depth = int(sys.argv[1]) if len(sys.argv) > 1 else 1
W, H = 1920, 1080
scene, cam = scenes.grid_scene(65536), rt.orbit_camera(3, 256, W, H)
cam.max_depth = float(depth)
with rt.Renderer(W, H) as r:
    r.set_scene(scene)
    r.init()
    for _ in range(4):
        r.update(cam)
    r.synchronize()
    print("rays", r.rays_traced(), "exact", r.exact_tests(), "tree", r.tree_info())
