"""Pinned inputs shared by the parity tests (SURVEY.md section 8(d) configurations)."""
import numpy as np

import gpu_ray_tracing_b200 as rt


def camera(width, height, max_depth=1, spp=1, random_seed=0.5, moved=1, defocus_angle=None, fov=None,
           look_from=None):
    s = rt.CameraSettings.default()
    s.max_depth = max_depth
    s.samples_per_pixel = spp
    s.camera_has_moved = moved
    if defocus_angle is not None:
        s.defocus_angle = defocus_angle
    if fov is not None:
        s.field_of_view = fov
    if look_from is not None:
        s.look_from = rt.RtVec3(*look_from)
    return rt.SceneCamera.from_settings(s, width, height, random_seed)


def default_scene():
    return rt.create_default_spheres(7)


def random_scene(n=1024):
    return rt.create_random_spheres(n, 1234)


def grid_scene(n=65536):
    return rt.create_grid_spheres(n, 4321)


def degenerate_scene():
    """A small scene with spheres that break every bound: NaN / infinite centres and radii, zero and negative radii,
    magnitudes beyond the 2^30 window, a huge ground sphere — next to ordinary ones."""
    rng = np.random.default_rng(11)
    n = 90
    sp = np.zeros(n, dtype=rt.SPHERE_DTYPE)
    sp["position"] = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    sp["position"][:, 1] = np.abs(sp["position"][:, 1]) * 0.2
    sp["radius"] = rng.uniform(0.1, 0.5, n).astype(np.float32)
    sp["color"] = rng.uniform(0.2, 1, (n, 4)).astype(np.float32)
    sp["color"][:, 3] = rng.choice([-2.0, 0.1, 2.0], n)
    sp["color"][sp["color"][:, 3] == 2.0, 0] = 1.5
    sp["position"][0] = (0, -1000, 0); sp["radius"][0] = 1000
    sp["radius"][7] = 0.0
    sp["radius"][13] = -0.4
    sp["position"][21] = (3e9, 0, 0); sp["radius"][21] = 2.9e9          # beyond the magnitude window, but in view
    sp["position"][33] = (np.inf, 0, 0)
    sp["radius"][47] = np.nan
    sp["position"][58] = (np.nan, 1, 1)
    sp["radius"][66] = np.inf
    sp["position"][80] = (1e30, 1e30, 1e30); sp["radius"][80] = 1e30    # the reference's own arithmetic overflows here
    return rt.SphereCollection(sp, n)
