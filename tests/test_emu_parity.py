"""CPU stand-in for the GPU parity tests: the kernels' own per-ray code (rt_core.cuh compiled for
the host) driven with the kernels' broad-phase / narrow-phase control flow must reproduce the
oracle bit for bit, and the broad phase must never reject a pair the exact quadratic accepts."""
import numpy as np
import pytest

import gpu_ray_tracing_b200 as rt
import oracle_lib as ol
import scenes

CASES = [
    ("default d30", lambda: scenes.default_scene(), 200, 112, dict(max_depth=30, spp=500)),
    ("default d1 pinhole", lambda: scenes.default_scene(), 256, 144, dict(max_depth=1, spp=1, defocus_angle=0.0)),
    ("random1024 d1", lambda: scenes.random_scene(1024), 160, 90, dict(max_depth=1, spp=1)),
    ("random1024 d12", lambda: scenes.random_scene(1024), 96, 54, dict(max_depth=12, spp=500)),
    ("grid8192 d4 low camera", lambda: scenes.grid_scene(8192), 96, 54, dict(max_depth=4, spp=500, look_from=(30.0, 1.5, 4.0), fov=50.0)),
    ("inside glass sphere", lambda: scenes.default_scene(), 64, 36, dict(max_depth=10, spp=500, look_from=(0.0, 1.0, 0.3))),
]


@pytest.mark.parametrize("variant", [0, 2], ids=["primary-pass", "general"])
@pytest.mark.parametrize("name,make,W,H,kw", CASES, ids=[c[0] for c in CASES])
def test_emulated_kernel_matches_oracle(name, make, W, H, kw, variant):
    scene, cam = make(), scenes.camera(W, H, **kw)
    cpu, rays = ol.oracle_update(cam, scene, W, H)
    emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=variant)
    if variant == 0:
        assert ol.EMU_STATS["primary_rays"] == W * H and ol.EMU_STATS["untraced"] == 0
    rep = ol.compare_images(cpu, emu)
    print(name, rep, "exact tests per ray: %.2f of %d spheres" % (exact / max(rays, 1), scene.count))
    assert violations == 0
    assert rays_e == rays
    assert rep["bit_identical"], rep
    assert exact < 0.08 * rays * scene.count + rays      # the broad phase actually culls


def test_emulated_progressive_frames_and_windows():
    W, H = 120, 68
    scene = scenes.default_scene()
    cams = [scenes.camera(W, H, max_depth=8, spp=500, random_seed=s, moved=m) for s, m in ((0.25, 1), (0.5, 0), (0.875, 0))]
    cpu = np.zeros((H, W, 4), np.float32)
    emu = np.zeros((H, W, 4), np.float32)
    for cam in cams:
        cpu, _ = ol.oracle_update(cam, scene, W, H, accum=cpu)
        emu, _, _, v = ol.emu_update(cam, scene, W, H, accum=emu)
        assert v == 0
    assert np.array_equal(cpu.view(np.uint32), emu.view(np.uint32)) and (cpu[..., 3] == 3).all()
    part, _, _, _ = ol.emu_update(cams[0], scene, W, H, row_begin=20, row_count=12)
    full, _ = ol.oracle_update(cams[0], scene, W, H)
    assert np.array_equal(part, full[20:32])


def test_broad_phase_is_conservative_on_adversarial_geometry():
    """Huge and tiny spheres, far-away origins, grazing rays: still no violation, still bit-exact."""
    rng = np.random.default_rng(5)
    n = 600
    sp = np.zeros(n, dtype=rt.SPHERE_DTYPE)
    sp["position"] = rng.uniform(-40, 40, (n, 3)).astype(np.float32)
    sp["radius"] = np.exp(rng.uniform(np.log(1e-3), np.log(30.0), n)).astype(np.float32)
    sp["position"][:4] = [(0, -1e5, 0), (0, 0, 5e3), (1e4, 1e4, 0), (0, 300, 0)]
    sp["radius"][:4] = [1e5, 4.9e3, 1e4, 299.5]
    sp["color"] = rng.uniform(0, 1, (n, 4)).astype(np.float32)
    sp["color"][:, 3] = rng.choice([-2.0, 0.1, 2.0], n)
    sp["color"][sp["color"][:, 3] == 2.0, 0] = 1.5
    scene = rt.SphereCollection(sp, n)
    for look_from in ((13.0, 2.0, 3.0), (900.0, 40.0, -700.0), (0.5, 0.01, 0.5)):
        W, H = 72, 40
        cam = scenes.camera(W, H, max_depth=6, spp=500, look_from=look_from, fov=60.0)
        cpu, rays = ol.oracle_update(cam, scene, W, H)
        for variant in (0, 2):
            emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=variant)
            assert violations == 0 and rays == rays_e
            assert np.array_equal(cpu.view(np.uint32), emu.view(np.uint32))
