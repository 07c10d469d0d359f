"""Small fixed workload for ncu launch lists: the app's own working point (default scene, 1280x720, max_depth 30)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

W, H = 1280, 720
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 4
scene, cam = scenes.default_scene(), scenes.camera(W, H, 30, 500)
with rt.Renderer(W, H) as r:
    r.set_scene(scene)
    r.init()
    r.update(cam)
    r.synchronize()
    r.timer_begin()
    for _ in range(iters):
        r.update(cam)
    print("ms per frame", r.timer_end() / iters)
    print("rays", r.rays_traced(), "per-bounce queue lengths", r.wavefront_counts()[:31])
