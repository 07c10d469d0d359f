#!/usr/bin/env python
"""Per-frame fixed cost of the depth-1 update: time per frame against the number of rows (1024-sphere scene), and the
linear fit t = a + b * pixels.  `a` is what a row shard pays however small it is (record prep, launches, ramp-up, tail)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

if __name__ == "__main__":
    W = int(sys.argv[1]) if len(sys.argv) > 1 else 7680
    Hfull = W * 9 // 16
    scene = scenes.random_scene(1024)
    xs, ys = [], []
    for rows in (8, 32, 128, 270, 540, 1080, 2160):
        if rows > Hfull:
            continue
        r0 = (Hfull - rows) // 2 // 8 * 8
        with rt.Renderer(W, Hfull, row_begin=r0, row_count=rows) as r:
            r.set_scene(scene)
            r.init()
            cam = scenes.camera(W, Hfull, 1, 1)
            for _ in range(5):
                r.update(cam)
            r.synchronize()
            r.timer_begin()
            n = 200 if rows <= 540 else 50
            for _ in range(n):
                r.update(cam)
            ms = r.timer_end() / n
        xs.append(rows * W / 1e6)
        ys.append(ms * 1e3)
        print(f"{W}x{rows:5d} rows (window at row {r0}): {ms * 1e3:8.1f} us/frame  {rows * W / ms / 1e3:9.0f} Mrays/s")
    b, a = np.polyfit(xs[2:], ys[2:], 1)
    print(f"fit over the larger windows: t = {a:.1f} us + {b:.2f} us per Mpixel")
