"""Developer probe (builds with -DRT_PRIM_PROBE=2 / 3): when the primary pass's warps end (2) or start (3), in ns after the
first warp's arrival, mean and max over the warps of one frame.  RT_B200_LIB selects the build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

if __name__ == "__main__":
    warps = 148 * 2 * 11
    out = []
    for name, W, H, bands in (("4K", 3840, 2160, None), ("8K 1/8", 7680, 4320, (16, 8, 0))):
        with rt.Renderer(W, H, bands=bands) as r:
            r.set_scene(scenes.random_scene(1024))
            r.init()
            cam = scenes.camera(W, H, 1, 1)
            for _ in range(5):
                r.update(cam)
            r.synchronize()
            means, maxs, mss = [], [], []
            for _ in range(5):
                r.reset_counters()
                r.synchronize()
                r.timer_begin()
                r.update(cam)
                mss.append(r.timer_end())
                means.append(r.exact_tests() / warps / 1e3)
                maxs.append(r.rays_traced() / 1e3)
            out.append(f"{name}: frame {min(mss) * 1e3:.1f} us, mean {sum(means) / 5:.1f} us, max {sum(maxs) / 5:.1f} us")
    print(os.path.basename(os.environ.get("RT_B200_LIB", "default")), " | ".join(out), flush=True)
