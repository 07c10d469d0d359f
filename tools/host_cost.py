"""Developer timing: host time per rt_update call (a tiny frame, so the GPU is never the limit) and the frame rate of
small depth-1 frames; RT_B200_LIB selects the build."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

if __name__ == "__main__":
    out = []
    for W, H in ((64, 36), (640, 360), (1280, 720), (1920, 1080)):
        with rt.Renderer(W, H) as r:
            r.set_scene(scenes.default_scene())
            r.init()
            cam = scenes.camera(W, H, 1, 1)
            for _ in range(20):
                r.update(cam)
            r.synchronize()
            n = 2000
            t0 = time.perf_counter()
            r.timer_begin()
            for _ in range(n):
                r.update(cam)
            t_host = (time.perf_counter() - t0) / n * 1e6
            ms = r.timer_end() / n
            out.append(f"{W}x{H}: enqueue {t_host:.1f} us/call, device {ms * 1e3:.1f} us/frame")
    print(os.path.basename(os.environ.get("RT_B200_LIB", "default")), " | ".join(out), flush=True)
