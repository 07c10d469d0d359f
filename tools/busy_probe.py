"""Developer probe (build with -DRT_PRIM_PROBE=1): how much of the primary pass's duration its warps are busy.  The kernel
reports the busy cycles summed over its warps through the exact-tests counter; mean busy time per warp / kernel time is
what the ramp, the ticket waits and the tail leave.  RT_B200_LIB selects the build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

if __name__ == "__main__":
    import torch
    mhz = float(os.environ.get("RT_PROBE_MHZ", "1965"))
    warps = 148 * 2 * 11
    out = []
    for name, W, H, bands in (("4K", 3840, 2160, None), ("8K 1/8", 7680, 4320, (16, 8, 0)), ("8K", 7680, 4320, None)):
        with rt.Renderer(W, H, bands=bands) as r:
            r.set_scene(scenes.random_scene(1024))
            r.init()
            cam = scenes.camera(W, H, 1, 1)
            for _ in range(5):
                r.update(cam)
            r.synchronize()
            r.reset_counters()
            n = 20
            r.timer_begin()
            for _ in range(n):
                r.update(cam)
            ms = r.timer_end() / n
            busy_ms = r.exact_tests() / n / warps / (mhz * 1e3)
            out.append(f"{name}: {ms:.4f} ms, mean busy {busy_ms:.4f} ms = {busy_ms / ms:.3f}")
    print(os.path.basename(os.environ.get("RT_B200_LIB", "default")), " | ".join(out), flush=True)
