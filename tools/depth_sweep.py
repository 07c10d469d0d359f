"""Developer timing: the default scene at 1280x720 for a range of max_depth (the slope is the cost of one late bounce);
RT_B200_LIB selects the build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

if __name__ == "__main__":
    W, H = 1280, 720
    out = []
    with rt.Renderer(W, H) as r:
        r.set_scene(scenes.default_scene())
        r.init()
        for depth in (3, 4, 5, 6, 8, 10, 15, 20, 30, 50):
            cam = scenes.camera(W, H, depth, 500)
            for _ in range(3):
                r.update(cam)
            r.synchronize()
            r.timer_begin()
            for _ in range(20):
                r.update(cam)
            out.append(f"d{depth}: {r.timer_end() / 20 * 1e3:.0f}")
    print(os.path.basename(os.environ.get("RT_B200_LIB", "default")), "us per frame", " | ".join(out), flush=True)
