"""Developer timing of one shard of the 8K strong-scaling frame (interleaved 16-row bands, shard 0 of N) without the gather;
RT_B200_LIB selects the build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402

if __name__ == "__main__":
    W, H = 7680, 4320
    scene = scenes.random_scene(1024)
    out = []
    for n in (1, 2, 4, 8):
        with rt.Renderer(W, H, bands=(16, n, 0)) as r:
            r.set_scene(scene)
            r.init()
            cams = [scenes.camera(W, H, 1, 1, random_seed=((f + 1) * 2654435761 % (1 << 24)) / float(1 << 24)) for f in range(40)]
            for c in cams[:5]:
                r.update(c)
            r.synchronize()
            r.timer_begin()
            for c in cams:
                r.update(c)
            out.append(f"1/{n}: {r.timer_end() / len(cams):.4f}")
    print(os.path.basename(os.environ.get("RT_B200_LIB", "default")), "8K shard ms", " | ".join(out), flush=True)
