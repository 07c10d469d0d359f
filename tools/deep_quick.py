import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import gpu_ray_tracing_b200 as rt, scenes
def t(scene, W, H, depth, n=10, prec=0):
    cam = scenes.camera(W, H, depth, 500)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene); r.init(); r.set_precision(prec)
        r.update(cam); r.synchronize(); r.reset_counters(); r.timer_begin()
        for _ in range(n): r.update(cam)
        ms = r.timer_end() / n
        return ms, r.rays_traced() / n, r.exact_tests() / n
for name, sc, W, H, d in (("r1024 4K d30", scenes.random_scene(1024), 3840, 2160, 30), ("default 720p d30", scenes.default_scene(), 1280, 720, 30),
                          ("default 4K d30", scenes.default_scene(), 3840, 2160, 30), ("r1024 4K d8", scenes.random_scene(1024), 3840, 2160, 8),
                          ("grid65536 1080p d8", scenes.grid_scene(65536), 1920, 1080, 8)):
    ms, rays, ex = t(sc, W, H, d)
    print(f"{name:22s} {ms:8.3f} ms  rays {rays:.0f}  exact/ray {ex / rays:.2f}")
