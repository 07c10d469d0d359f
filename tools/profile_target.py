"""Small fixed workload for ncu: the headline configuration (1024 spheres, 3840x2160, depth 1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

variant = int(sys.argv[1]) if len(sys.argv) > 1 else 0
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 4
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 1
W, H = 3840, 2160
scene, cam = scenes.random_scene(1024), scenes.camera(W, H, depth, 1)
with rt.Renderer(W, H) as r:
    r.set_kernel_variant(variant)
    r.set_scene(scene)
    r.init()
    for _ in range(iters):
        r.update(cam)
    r.synchronize()
    print("rays", r.rays_traced())
    if depth > 1:
        print("per-bounce queue lengths", r.wavefront_counts()[:depth + 1])
