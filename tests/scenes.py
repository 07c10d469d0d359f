"""Pinned inputs shared by the parity tests (SURVEY.md section 8(d) configurations)."""
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
