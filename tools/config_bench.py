"""BASELINE.json configs 4 and 5 at full size; one process drives all the GPUs of the box.

  config 4: 1024 random spheres at 7680x4320, one image sharded by row blocks over N devices (rt_multi_*),
            RGBA8 gathered into devices[0] by peer copies over NVLink; reported with and without the gather.
  config 5: 65,536 spheres, 256-pose orbit fly-through at 1920x1080, frames dealt round-robin over N devices,
            every frame read back to pinned host memory (rt_read_rgba8_async).

Usage: python tools/config_bench.py [--gpus 1,2,4,8] [--out profiles/r1_configs45.json]
Timing is wall clock around fully synchronised regions (several devices, so no single stream's events apply);
each figure is the best of `--reps` repetitions after one warm-up.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpu_ray_tracing_b200 as rt  # noqa: E402
import scenes  # noqa: E402


def config4(n_dev, reps):
    W, H, N = 7680, 4320, 1024
    scene, cam = scenes.random_scene(N), scenes.camera(W, H, 1, 1)
    with rt.MultiRenderer(list(range(n_dev)), W, H) as m:
        m.set_scene(scene)
        m.init()
        best = {"update": 1e9, "update+gather": 1e9}
        for rep in range(reps + 1):
            m.gather_rgba8(to_host=False)            # drains every shard's stream
            t0 = time.perf_counter()
            m.update(cam)
            m.gather_rgba8(to_host=False)
            t1 = time.perf_counter()
            # kernels only: a second update, synchronised through the (cheap) ray counters
            m.rays_traced()
            t2 = time.perf_counter()
            m.update(cam)
            m.rays_traced()
            t3 = time.perf_counter()
            t4 = time.perf_counter()
            m.rays_traced()
            sync_cost = time.perf_counter() - t4
            if rep:
                best["update+gather"] = min(best["update+gather"], t1 - t0)
                best["update"] = min(best["update"], t3 - t2 - sync_cost)
        rays = W * H
        img = m.gather_rgba8()
    return {"config": 4, "n_gpus": n_dev, "width": W, "height": H, "spheres": N,
            "ms_update": best["update"] * 1e3, "ms_update_gather": best["update+gather"] * 1e3,
            "mrays_s_update": rays / best["update"] / 1e6, "mrays_s_update_gather": rays / best["update+gather"] / 1e6,
            "checksum": int(img[::97, ::89].astype(np.int64).sum())}


def config5(n_dev, reps):
    import torch
    W, H, N, POSES = 1920, 1080, 65536, 256
    scene = scenes.grid_scene(N)
    cams = [rt.orbit_camera(k, POSES, W, H) for k in range(POSES)]
    renderers = [rt.Renderer(W, H, device=d) for d in range(n_dev)]
    frames = torch.empty((POSES, H, W, 4), dtype=torch.uint8).pin_memory()
    try:
        for r in renderers:
            r.set_scene(scene)
            r.init()
            r.set_rgba8_output(True)     # the frame that is read back comes out of the shading pass
        best = 1e9
        for rep in range(reps + 1):
            poses = range(POSES) if rep else range(2 * n_dev)     # short warm-up
            t0 = time.perf_counter()
            for k in poses:
                r = renderers[k % n_dev]
                r.update(cams[k])                                   # camera_has_moved = 1: each pose starts a fresh image
                r.read_rgba8_async_into(frames[k].data_ptr())
            for r in renderers:
                r.read_wait()
            dt = time.perf_counter() - t0
            if rep:
                best = min(best, dt)
        rays = sum(r.rays_traced() for r in renderers)
    finally:
        for r in renderers:
            r.close()
    return {"config": 5, "n_gpus": n_dev, "width": W, "height": H, "spheres": N, "poses": POSES,
            "seconds": best, "mrays_s": POSES * W * H / best / 1e6, "ttests_s": POSES * W * H * N / best / 1e12,
            "checksum": int(frames[::17, ::53, ::47].to(torch.int64).sum().item()), "rays_counted_all_reps": int(rays)}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--configs", default="4,5")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    have = rt.device_count()
    results = []
    for n in [int(x) for x in args.gpus.split(",")]:
        if n > have:
            print(f"skipping N={n}: {have} device(s)", file=sys.stderr)
            continue
        if "4" in args.configs.split(","):
            results.append(config4(n, args.reps))
            print(json.dumps(results[-1]), flush=True)
        if "5" in args.configs.split(","):
            results.append(config5(n, 1 if n == 1 else args.reps))
            print(json.dumps(results[-1]), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)
