"""rt_update_static — the entry point that allocates nothing (scratch in a file-scope device array; no counterpart in
the reference).  Called through the C ABI with device pointers and checked
against the existing entry point (rt_update on a renderer handle) on two calls in a row: the bar is the one every other
GPU test here uses, bit-identical rgba32f accumulators.  In a file of its own that pytest collects last;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import gpu_ray_tracing_b200 as rt
import scenes

pytestmark = pytest.mark.gpu


def _reference_frames(scene, cams, W, H):
    """The existing entry point: init, then rt_update per camera; the accumulation image after each."""
    out = []
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        for cam in cams:
            r.update(cam)
            out.append(r.read_accum().copy())
    return out


def _free_bytes():
    from cuda.bindings import runtime as cudart
    err, free, _total = cudart.cudaMemGetInfo()
    assert int(err) == 0
    return int(free)


@pytest.mark.parametrize("n_spheres,depth", [(1024, 1), (199, 6)])
def test_static_update_matches_rt_update_twice_in_a_row(n_spheres, depth):
    import torch
    W, H = 640, 360
    scene = scenes.random_scene(n_spheres) if n_spheres == 1024 else scenes.default_scene()
    # two progressive frames: the second call reads the first one's texels (camera_has_moved = 0) and a re-rolled seed
    cams = [scenes.camera(W, H, depth, 8, random_seed=s, moved=m) for s, m in ((0.375, 1), (0.8125, 0))]
    want = _reference_frames(scene, cams, W, H)

    dev = torch.device("cuda", 0)
    spheres = torch.from_numpy(scene.spheres.view(np.uint8).reshape(-1).copy()).to(dev)     # the caller's input, on the device
    accum = torch.zeros((H, W, 4), dtype=torch.float32, device=dev)                          # input / output (`init` = zeros)
    torch.cuda.synchronize()
    for k, cam in enumerate(cams):
        before = _free_bytes() if k == 1 else None
        rt.update_static(0, spheres.data_ptr(), scene.count, cam, W, H, accum.data_ptr(), 0)
        torch.cuda.synchronize()
        if before is not None:
            assert _free_bytes() == before          # the second call (set-up done) changed nothing in device memory
        got = accum.cpu().numpy()
        assert np.array_equal(got.view(np.uint32), want[k].view(np.uint32)), (n_spheres, depth, k)
    assert (accum.cpu().numpy()[..., 3] == 2.0).all()


def test_static_update_rejects_what_it_cannot_hold():
    import torch
    lib = rt.lib()
    cam = scenes.camera(64, 36, 1, 1)
    accum = torch.zeros((36, 64, 4), dtype=torch.float32, device="cuda")
    spheres = torch.zeros(32 * (rt.RT_STATIC_MAX_SPHERES + 1), dtype=torch.uint8, device="cuda")
    rc = lib.rt_update_static(0, C.c_void_p(spheres.data_ptr()), rt.RT_STATIC_MAX_SPHERES + 1, C.byref(cam), 64, 36,
                              C.c_void_p(accum.data_ptr()), None)
    assert rc == rt.api.RT_ERR_INVALID_ARG and b"4096" in lib.rt_last_error()
    assert lib.rt_update_static(0, None, 5, C.byref(cam), 64, 36, C.c_void_p(accum.data_ptr()), None) == rt.api.RT_ERR_INVALID_ARG
    assert lib.rt_update_static(0, C.c_void_p(spheres.data_ptr()), 5, C.byref(cam), 64, 36, None, None) == rt.api.RT_ERR_INVALID_ARG
    # the maximum itself is accepted (zero-radius spheres at the origin: nothing but sky)
    rt.update_static(0, spheres.data_ptr(), rt.RT_STATIC_MAX_SPHERES, cam, 64, 36, accum.data_ptr(), 0)
    torch.cuda.synchronize()
    assert (accum.cpu().numpy()[..., 3] == 1.0).all()
