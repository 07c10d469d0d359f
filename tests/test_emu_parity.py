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
        assert ol.EMU_STATS["tile_dropped"] == 0       # the tile level never removes a record a ray test keeps
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


def test_nearest_first_cull_on_grazing_chains():
    """The "already beaten" cull of the primary pass (rtc::primary_near): chains of overlapping spheres strung along
    the view direction, near and 10,000 units away, seen through narrow fields of view so that most hits are grazing
    ones — where the reference's own root is least accurate.  The emulation checks the bound against every root the
    reference accepts and every skipped pair against the final hit: 0 violations, image bit-identical, and the cull
    does remove tests."""
    rng = np.random.default_rng(77)
    culled = 0
    for dist, radius, fov in ((30.0, 1.0, 8.0), (800.0, 6.0, 2.0), (10000.0, 40.0, 1.0), (10000.0, 3.0, 0.1)):
        n = 160
        sp = np.zeros(n, dtype=rt.SPHERE_DTYPE)
        t = np.sort(rng.uniform(0.3, 1.7, n)) * dist
        axis = np.array([1.0, 0.1, -0.4]) / np.linalg.norm([1.0, 0.1, -0.4])
        side = rng.normal(0.0, 0.9 * radius, (n, 3))
        sp["position"] = (t[:, None] * axis[None, :] + side).astype(np.float32)
        sp["radius"] = (radius * rng.uniform(0.5, 1.5, n)).astype(np.float32)
        sp["color"] = rng.uniform(0.1, 1, (n, 4)).astype(np.float32)
        sp["color"][:, 3] = rng.choice([-2.0, 0.2, 2.0], n)
        sp["color"][sp["color"][:, 3] == 2.0, 0] = 1.5
        scene = rt.SphereCollection(sp, n)
        W, H = 64, 32
        for defocus in (0.0, 0.6):
            s = rt.CameraSettings.default()
            s.max_depth, s.samples_per_pixel, s.camera_has_moved = 1, 1, 1
            s.look_from = rt.RtVec3(0.0, 0.0, 0.0)
            s.look_at = rt.RtVec3(*(float(v) for v in axis * dist))
            s.field_of_view, s.defocus_angle, s.focus_distance = fov, defocus, dist
            cam = rt.SceneCamera.from_settings(s, W, H, 0.25)
            cpu, rays = ol.oracle_update(cam, scene, W, H)
            emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=0)
            assert violations == 0 and rays == rays_e, (dist, defocus, violations)
            assert np.array_equal(cpu.view(np.uint32), emu.view(np.uint32)), (dist, defocus)
            culled += ol.EMU_STATS["beaten"]
    assert culled > 0


def test_randomised_scenes_and_cameras_audit():
    """Seeded random scenes (log-uniform radii over 5 decades, clustered and overlapping spheres) seen by random
    cameras — pinhole and wide defocus disks, inside clusters, far away, huge and tiny fields of view: neither broad
    phase may reject a pair the reference hits, and the image stays bit-identical."""
    rng = np.random.default_rng(20261104)
    for case in range(24):
        n = int(rng.integers(1, 400))
        sp = np.zeros(n, dtype=rt.SPHERE_DTYPE)
        spread = float(np.exp(rng.uniform(np.log(0.5), np.log(500.0))))
        sp["position"] = rng.normal(0.0, spread, (n, 3)).astype(np.float32)
        sp["radius"] = np.exp(rng.uniform(np.log(1e-3), np.log(1e2), n)).astype(np.float32) * np.float32(spread / 10.0)
        sp["color"] = rng.uniform(0, 1, (n, 4)).astype(np.float32)
        sp["color"][:, 3] = rng.choice([-2.0, 0.0, 0.3, 2.0], n)
        sp["color"][sp["color"][:, 3] == 2.0, 0] = 1.5
        scene = rt.SphereCollection(sp, n)
        look_from = tuple(float(v) for v in rng.normal(0.0, spread * float(rng.choice([0.3, 1.0, 4.0, 40.0])), 3))
        W, H = 40, 22
        cam = scenes.camera(W, H, max_depth=int(rng.integers(1, 6)), spp=500, look_from=look_from,
                            fov=float(rng.choice([1.0, 20.0, 90.0, 150.0])),
                            defocus_angle=float(rng.choice([0.0, 0.6, 10.0, 60.0])), random_seed=float(rng.uniform(0, 1)))
        cpu, rays = ol.oracle_update(cam, scene, W, H)
        for variant in (0, 2):
            emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=variant)
            assert violations == 0, (case, variant, violations)
            if variant == 0:
                assert ol.EMU_STATS["tile_dropped"] == 0, (case, ol.EMU_STATS)
            assert rays_e == rays
            assert np.array_equal(cpu.view(np.uint32), emu.view(np.uint32)), (case, variant)


def test_tile_level_culls_and_changes_nothing():
    """The per-tile test of the primary pass (32 x 8 pixel tiles): it keeps a small part of the (tile, record) pairs
    at a resolution where tiles are narrow, never drops a record one of the tile's rays would test, and the exact-test
    count is the per-ray broad phase's own."""
    W, H = 640, 360
    scene = scenes.random_scene(1024)
    for defocus in (None, 0.0):
        cam = scenes.camera(W, H, max_depth=1, spp=1, defocus_angle=defocus)
        cpu, rays = ol.oracle_update(cam, scene, W, H)
        emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=0)
        st = dict(ol.EMU_STATS)
        n_tiles = ((W + 31) // 32) * ((H + 7) // 8)
        kept = st["tile_pairs"] / (n_tiles * scene.count)
        print("defocus", defocus, "tile level keeps %.2f %% of the pairs, %.1f records per tile" % (100 * kept, kept * scene.count), st)
        assert violations == 0 and st["tile_dropped"] == 0 and rays_e == rays == W * H
        assert np.array_equal(cpu.view(np.uint32), emu.view(np.uint32))
        assert kept < 0.10


def test_primary_records_are_conservative_by_construction():
    """primary_rho / primary_record invariants checked directly: always-candidate record when the camera centre is
    inside or within the slack of the (inflated) sphere, a negative threshold otherwise, and rho >= the largest
    |origin - centre| the shader itself produces for that camera."""
    import ctypes as C
    lib = ol.emu()
    lib.emu_primary_record.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.emu_primary_rho.restype = C.c_double
    lib.emu_primary_rho.argtypes = [C.c_void_p]
    W, H = 64, 36
    for defocus in (0.0, 0.6, 25.0):
        cam = scenes.camera(W, H, defocus_angle=defocus)
        rho = lib.emu_primary_rho(C.byref(cam))
        assert rho >= 0.0 and (rho == 0.0) == (defocus == 0.0)
        # origins the shader generates (oracle_get_ray) never leave the rho ball
        out = np.zeros(6, dtype=np.float32)
        worst = 0.0
        for s in range(0, 4000, 7):
            ol.oracle().oracle_get_ray(C.byref(cam), s % W, (s // W) % H, 12345 + s, out.ctypes.data_as(C.c_void_p))
            worst = max(worst, float(np.linalg.norm(out[:3].astype(np.float64) - np.array(cam.center.tuple(), dtype=np.float64))))
        assert worst <= rho
        sp = np.zeros(3, dtype=rt.SPHERE_DTYPE)
        c = np.array(cam.center.tuple(), dtype=np.float32)
        sp["position"][0], sp["radius"][0] = c + np.float32(0.5), 2.0          # centre inside
        sp["position"][1], sp["radius"][1] = c + np.array([3.0, 0, 0], np.float32), 3.0 - 1e-5   # just outside, within the slack
        sp["position"][2], sp["radius"][2] = c + np.array([0, 0, -50.0], np.float32), 1.0        # well outside
        rec = np.zeros((3, 4), dtype=np.float32)
        for i in range(3):
            lib.emu_primary_record(C.byref(cam), sp[i:i + 1].ctypes.data_as(C.c_void_p), rec[i].ctypes.data_as(C.c_void_p))
        assert tuple(rec[0]) == (0.0, 0.0, 0.0, 1.0) and tuple(rec[1]) == (0.0, 0.0, 0.0, 1.0)
        assert rec[2][3] < 0.0 and abs(-rec[2][3] - np.sqrt(50.0 ** 2 - (1.0 + rho) ** 2)) < 0.02


def test_degenerate_spheres_follow_the_reference_loop():
    """NaN roots make sphere_list_hit order dependent; the tree traversal must notice and retrace such rays in index
    order, and no bound may hide a sphere whose reference arithmetic goes non-finite."""
    scene = scenes.degenerate_scene()
    W, H = 64, 36
    for depth in (1, 5):
        cam = scenes.camera(W, H, max_depth=depth, spp=500, look_from=(9.0, 1.5, 2.0), fov=50.0)
        cpu, rays = ol.oracle_update(cam, scene, W, H)
        for variant in (0, 2):
            emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=variant)
            assert violations == 0 and rays == rays_e, (depth, variant, violations)
            assert ol.nan_aware_equal(cpu, emu), (depth, variant)
    cam = scenes.camera(W, H, max_depth=5, spp=500, look_from=(9.0, 1.5, 2.0), fov=50.0)
    ol.emu_update(cam, scene, W, H, variant=0)
    assert ol.EMU_STATS["cluster_rays"] > 0 and ol.EMU_STATS["retraced"] > 0


def test_tree_culls_bounce_rays():
    """The bounding-sphere tree: a bounce ray of the 1024-sphere scene tests a small part of the scene, no ray is
    retraced, nothing is missed."""
    W, H = 128, 72
    scene = scenes.random_scene(1024)
    cam = scenes.camera(W, H, max_depth=8, spp=500)
    cpu, rays = ol.oracle_update(cam, scene, W, H)
    emu, rays_e, exact, violations = ol.emu_update(cam, scene, W, H, variant=0)
    st = dict(ol.EMU_STATS)
    per_ray = (st["cluster_tests"] + st["member_tests"]) / st["cluster_rays"]
    print("tree: %.1f record tests per bounce ray of %d spheres" % (per_ray, scene.count), st)
    assert violations == 0 and rays == rays_e and st["retraced"] == 0
    assert np.array_equal(cpu.view(np.uint32), emu.view(np.uint32))
    assert per_ray < 0.3 * scene.count
