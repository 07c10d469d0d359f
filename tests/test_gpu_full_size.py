"""GPU parity at the sizes BASELINE.json names, the exports round 1 left without a test of their own, the tolerance
mode, and the fused gather.  Everything goes through the C ABI; the checker is the CPU oracle on the same bytes.

  config 3  1024 random spheres, 3840x2160      whole frame (also compared in every bench.py run)
  config 4  1024 random spheres, 7680x4320      whole frame, single renderer and row-sharded (rt_multi_*)
  config 5  65,536 grid spheres, 1920x1080      one orbit pose at depth 1; 480x270 at depth 8 (tree walked in global memory)
"""
import numpy as np
import pytest

import gpu_ray_tracing_b200 as rt
import oracle_lib as ol
import scenes

pytestmark = pytest.mark.gpu


def assert_bit_identical(gpu, cpu, what):
    rep = ol.compare_images(cpu, gpu)
    print(what, rep)
    assert rep["pixels_over_1_255"] == 0 and rep["bit_identical"], (what, rep)


def device_bytes(ptr, nbytes):
    import torch
    from cuda.bindings import runtime as cudart
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    err, = cudart.cudaMemcpy(host.data_ptr(), ptr, nbytes, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
    assert int(err) == 0
    return host.numpy().copy()


def device_fill(ptr, value, nbytes):
    from cuda.bindings import runtime as cudart
    err, = cudart.cudaMemset(ptr, value, nbytes)
    assert int(err) == 0


# ------------------------------------------------------------------------------------------------
# BASELINE sizes
# ------------------------------------------------------------------------------------------------
def test_config3_whole_4k_frame():
    """The headline frame, all 8,294,400 texels."""
    W, H = 3840, 2160
    scene, cam = scenes.random_scene(1024), scenes.camera(W, H, 1, 1, random_seed=0.3125)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        gpu, rays = r.read_accum(), r.rays_traced()
    assert rays == rays_cpu == W * H
    assert_bit_identical(gpu, cpu, "config 3, 3840x2160 whole frame")


def test_config4_8k_frame_single_and_row_sharded():
    """7680x4320 / 1024 spheres / depth 1: the oracle frame against the single-renderer image AND the row-sharded
    one (rt_multi_*, as many distinct devices as the box has; gathered by the fused stores into devices[0])."""
    W, H = 7680, 4320
    scene, cam = scenes.random_scene(1024), scenes.camera(W, H, 1, 1, random_seed=0.6875)
    cpu, _ = ol.oracle_update(cam, scene, W, H)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        gpu = r.read_accum()
    assert_bit_identical(gpu, cpu, "config 4, 7680x4320 single renderer")
    del gpu
    have = rt.device_count()
    n_shards = max(2, min(have, 8))
    with rt.MultiRenderer([i % have for i in range(n_shards)], W, H) as m:
        m.set_scene(scene)
        m.init()
        m.update(cam)
        rgba = m.gather_rgba8()
        accum = m.gather_accum()
        assert m.rays_traced() == W * H
    assert_bit_identical(accum, cpu, f"config 4, {n_shards} row shards over {min(have, n_shards)} device(s)")
    assert np.array_equal(rgba, ol.quantize(cpu))


def test_config5_65536_spheres_1080p_and_depth8():
    """65,536 spheres (seed 4321): one orbit pose at 1920x1080, depth 1 — the hot loop of sphere_list_hit
    (compute_shader.wgsl:164-180) over the whole sphere buffer — and 480x270 at depth 8 (bounce rays walk the tree in
    global memory)."""
    scene = scenes.grid_scene(65536)
    W, H = 1920, 1080
    cam = rt.orbit_camera(37, 256, W, H)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        gpu, rays = r.read_accum(), r.rays_traced()
        assert r.tree_info()[2] == 2 and r.tree_build_ms() > 0.0
    assert rays == rays_cpu == W * H
    assert_bit_identical(gpu, cpu, "config 5, 65,536 spheres 1920x1080 depth 1")
    W, H = 480, 270
    cam = rt.orbit_camera(101, 256, W, H)
    cam.max_depth = 8.0
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        gpu, rays = r.read_accum(), r.rays_traced()
    assert rays == rays_cpu
    assert_bit_identical(gpu, cpu, "config 5 scene, 480x270 depth 8")


# ------------------------------------------------------------------------------------------------
# mismatch count against a differently rounded implementation (the tolerance the port target states)
# ------------------------------------------------------------------------------------------------
def test_gpu_vs_libm_oracle_mismatch_counts():
    """GPU (exact mode) vs liboracle_libm.so — the restatement with libm sinf / cosf / powf in place of the pinned
    polynomials, standing in for a WGSL implementation whose builtins round differently (compute_shader.wgsl:140,
    239-240, 329).  Tolerance: <= 1/255 per channel after RGBA8 quantisation; the count of pixels over it is reported
    and bounded.  Depth 1 differs by colour rounding and the odd material-branch flip; depth 30 is a chaotic path."""
    W, H = 1280, 720
    scene = scenes.random_scene(1024)
    for depth, bound in ((1, 2e-5), (30, 0.05)):
        cam = scenes.camera(W, H, depth, 500)
        libm, _ = ol.oracle_update(cam, scene, W, H, libm=True)
        with rt.Renderer(W, H) as r:
            r.set_scene(scene)
            r.init()
            r.update(cam)
            gpu = r.read_accum()
        rep = ol.compare_images(libm, gpu)
        print(f"GPU vs libm oracle, depth {depth}:", rep)
        assert rep["pixels_over_1_255"] <= bound * rep["pixels"], (depth, rep)


@pytest.mark.parametrize("level", [rt.RT_PRECISION_FAST_SHADING, rt.RT_PRECISION_FAST_ALL])
@pytest.mark.parametrize("depth", [1, 8])
def test_fast_precision_mode_within_tolerance(depth, level):
    """rt_set_precision: FAST_SHADING keeps the primary rays and every hit decision of the exact mode, so a depth-1
    image may differ by colour rounding and the odd material-branch decision only; FAST_ALL also approximates get_ray,
    which moves samples across silhouettes (a few pixels per 100,000).  Counts are reported and bounded."""
    W, H = 1280, 720
    scene, cam = scenes.random_scene(1024), scenes.camera(W, H, depth, 500)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.set_precision(level)
        r.update(cam)
        fast, rays_fast = r.read_accum(), r.rays_traced()
        r.set_precision(rt.RT_PRECISION_EXACT)
        r.reset_counters()
        r.update(cam)
        exact = r.read_accum()
    rep = ol.compare_images(cpu, fast)
    print(f"precision level {level} vs oracle, depth {depth}:", rep, "rays", rays_fast, rays_cpu)
    assert_bit_identical(exact, cpu, "exact mode after switching back")
    assert (fast[..., 3] == 1.0).all()
    if depth == 1:
        assert rays_fast == rays_cpu
        assert rep["pixels_over_1_255"] <= (4 if level == rt.RT_PRECISION_FAST_SHADING else 1e-4 * rep["pixels"]), rep
    else:
        assert abs(rays_fast - rays_cpu) <= 0.01 * rays_cpu
        assert rep["pixels_over_1_255"] <= 0.05 * rep["pixels"], rep
    with rt.Renderer(8, 8) as r2, pytest.raises(rt.RtError):
        r2.set_precision(7)


# ------------------------------------------------------------------------------------------------
# exports without a test of their own in round 1
# ------------------------------------------------------------------------------------------------
def test_update_n_equals_n_updates():
    """rt_update_n(cameras, n) == n x rt_update: progressive frames with per-frame random_seed (camera.rs:346), the
    reset flag on the first only; depth 1 (fused pass) and depth 6 (wavefront)."""
    W, H = 333, 187
    scene = scenes.default_scene()
    for depth in (1, 6):
        cams = [scenes.camera(W, H, depth, 500, random_seed=(i * 37 % 64) / 64.0, moved=1 if i == 0 else 0) for i in range(9)]
        with rt.Renderer(W, H) as r:
            r.set_scene(scene)
            r.init()
            for c in cams:
                r.update(c)
            one_by_one, rays_a = r.read_accum(), r.rays_traced()
            r.init()
            r.reset_counters()
            r.update_n(cams)
            batched, rays_b = r.read_accum(), r.rays_traced()
        assert rays_a == rays_b
        assert np.array_equal(one_by_one.view(np.uint32), batched.view(np.uint32))
        assert (batched[..., 3] == 9.0).all()
        cpu = None
        for c in cams:
            cpu, _ = ol.oracle_update(c, scene, W, H, accum=cpu)
        assert_bit_identical(batched, cpu, f"rt_update_n, 9 frames, depth {depth}")


def test_render_frames_matches_per_frame_renders():
    """rt_render_frames: one RGBA8 frame per camera == init + update + read for each."""
    W, H = 200, 120
    scene = scenes.random_scene(300)
    cams = [rt.orbit_camera(k, 5, W, H, radius=14.0, cam_height=3.0, fov_deg=30.0) for k in range(5)]
    frames = rt.render_frames(scene, cams, W, H)
    assert frames.shape == (5, H, W, 4)
    for k, cam in enumerate(cams):
        cpu, _ = ol.oracle_update(cam, scene, W, H)
        assert np.array_equal(frames[k], ol.quantize(cpu)), k


def test_quantize_into_device_buffer_and_interleaved_reads():
    """rt_quantize_into (caller-owned device buffer) equals the oracle's quantisation; a blocking rt_read_rgba8 issued
    while an asynchronous read of the previous frame may still be copying does not tear either frame."""
    import torch
    W, H = 640, 360
    scene = scenes.default_scene()
    nbytes = W * H * 4
    dst = rt.device_alloc(0, nbytes)
    try:
        with rt.Renderer(W, H) as r:
            r.set_scene(scene)
            r.init()
            cam_a, cam_b = scenes.camera(W, H, 4, 8, random_seed=0.25), scenes.camera(W, H, 4, 8, random_seed=0.75)
            r.update(cam_a)
            r.quantize_into(dst)
            r.synchronize()
            want_a = ol.quantize(ol.oracle_update(cam_a, scene, W, H)[0])
            assert np.array_equal(device_bytes(dst, nbytes).reshape(H, W, 4), want_a)
            host_a = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
            for _ in range(4):
                r.update(cam_a)
                r.read_rgba8_async_into(host_a.data_ptr())      # slot 0 or 1 in flight
                r.update(cam_b)
                got_b = r.read_rgba8()                          # blocking read right behind it
                r.read_wait()
                assert np.array_equal(host_a.numpy(), want_a)
                assert np.array_equal(got_b, ol.quantize(ol.oracle_update(cam_b, scene, W, H)[0]))
    finally:
        rt.device_free(0, dst)


def test_fused_rgba8_target_and_stream_flags():
    """rt_set_rgba8_target: the shading pass stores the window's RGBA8 texels at a caller-given device address (here a
    row window into the middle of a full-frame buffer, as a shard of a sharded image does); rt_stream_signal /
    rt_stream_wait order a consumer behind them, and a wait nobody answers gives up after its timeout instead of
    hanging the device."""
    W, H, r0, rows = 512, 300, 96, 120
    scene, cam = scenes.random_scene(200), scenes.camera(W, H, 1, 1)
    full = rt.device_alloc(0, W * H * 4 + 256)
    flags = full + W * H * 4
    try:
        device_fill(full, 0x5A, W * H * 4)
        device_fill(flags, 0, 256)
        with rt.Renderer(W, H, row_begin=r0, row_count=rows) as r:
            r.set_scene(scene)
            r.init()
            r.set_rgba8_target(full)                             # texel (0, 0) of the FULL image
            before = r.kernel_launches()
            r.update(cam)
            assert r.kernel_launches() - before == 2            # records + primary pass: no quantise, no copy
            r.stream_signal(flags, 7)
            r.stream_wait(flags, 1, 7, 1000)
            r.synchronize()
            assert r.stream_wait_timed_out() == 0
            accum = r.read_accum()
            img = device_bytes(full, W * H * 4).reshape(H, W, 4)
            assert np.array_equal(img[r0:r0 + rows], ol.quantize(accum))
            assert (img[:r0] == 0x5A).all() and (img[r0 + rows:] == 0x5A).all()
            cpu, _ = ol.oracle_update(cam, scene, W, H, row_begin=r0, row_count=rows)
            assert np.array_equal(img[r0:r0 + rows], ol.quantize(cpu))
            assert device_bytes(flags, 4).view(np.uint32)[0] == 7
            # a deep frame does not use the target (the texels are finished by the bounce kernels): reads still work
            r.set_rgba8_target(None)
            r.stream_wait(flags + 4, 2, 1, 30)                  # two flags nobody sets
            r.synchronize()
            assert r.stream_wait_timed_out() == 2
    finally:
        rt.device_free(0, full)


@pytest.mark.skipif(rt.device_count() < 2, reason="needs two devices (real peer stores over NVLink)")
def test_multi_renderer_on_distinct_devices():
    """rt_multi_* with one shard per device: the fused stores into devices[0] cross NVLink; bit-identical to one GPU."""
    W, H = 1920, 1080
    n = min(rt.device_count(), 8)
    scene = scenes.random_scene(1024)
    for kw in (dict(max_depth=1, spp=1), dict(max_depth=5, spp=3)):
        cam = scenes.camera(W, H, **kw)
        with rt.Renderer(W, H) as r:
            r.set_scene(scene)
            r.init()
            r.update(cam)
            single = r.read_accum()
        with rt.MultiRenderer(list(range(n)), W, H) as m:
            m.set_scene(scene)
            m.init()
            m.update(cam)
            rgba = m.gather_rgba8()
            accum = m.gather_accum()
        assert np.array_equal(accum.view(np.uint32), single.view(np.uint32))
        assert np.array_equal(rgba, ol.quantize(single))


@pytest.mark.parametrize("depth", [1, 6])
def test_interleaved_band_shards_equal_the_whole_image(depth):
    """rt_create_bands: three shards owning every third 8-row band (ragged height: the last band is short) render exactly
    the texels of the single renderer; at depth 1 their fused stores assemble the RGBA8 image in one full-frame buffer."""
    W, H, band, n = 500, 203, 8, 3
    scene, cam = scenes.random_scene(400), scenes.camera(W, H, depth, 4)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        whole, rays_whole = r.read_accum(), r.rays_traced()
    full = rt.device_alloc(0, 2 * W * H * 4 + 256)
    full2 = full + W * H * 4
    try:
        device_fill(full, 0, 2 * W * H * 4 + 256)
        got = np.zeros_like(whole)
        rays = 0
        rows_seen = []
        for idx in range(n):
            with rt.Renderer(W, H, bands=(band, n, idx)) as r:
                assert r.row_count == rt.shard_bands(H, band, n, idx)
                r.set_scene(scene)
                r.init()
                r.set_rgba8_target(full)
                r.update(cam)
                rows = r.image_rows()
                got[rows] = r.read_accum()
                rays += r.rays_traced()
                rows_seen.append(rows)
                # the pipelined gather: the same rows scattered into a second full image by a copy engine, flag behind
                r.send_rgba8_async(full2)
                r.copy_signal(full2 + W * H * 4, idx + 1)
                r.copy_wait(full2 + W * H * 4, 1, idx + 1, 1000)
                r.read_wait()
                assert r.stream_wait_timed_out() == 0
        assert sorted(np.concatenate(rows_seen).tolist()) == list(range(H))
        assert rays == rays_whole
        assert np.array_equal(got.view(np.uint32), whole.view(np.uint32))
        if depth == 1:
            assert np.array_equal(device_bytes(full, W * H * 4).reshape(H, W, 4), ol.quantize(whole))
        assert np.array_equal(device_bytes(full2, W * H * 4).reshape(H, W, 4), ol.quantize(whole))
    finally:
        rt.device_free(0, full)
    have = rt.device_count()
    with rt.MultiRenderer([i % have for i in range(n)], W, H, band_rows=band) as m:
        m.set_scene(scene)
        m.init()
        m.update(cam)
        assert np.array_equal(m.gather_rgba8(), ol.quantize(whole))
        assert np.array_equal(m.gather_accum().view(np.uint32), whole.view(np.uint32))


def test_converged_default_image_500_spp():
    """The application's working point (camera.rs:33-34: 500 samples per pixel, max_depth 30; default scene) at reduced
    size: 500 progressive frames through rt_update_n — random_seed re-rolled per frame (camera.rs:346), reset flag on the
    first only — against the oracle looped 500 times.  Every texel holds n = 500 and the running means agree bit for bit
    (the running-mean order is the frames' order on both sides)."""
    W, H, S = 128, 72, 500
    scene = scenes.default_scene()
    cams = [scenes.camera(W, H, 30, S, random_seed=((i + 1) * 2654435761 % (1 << 24)) / float(1 << 24), moved=1 if i == 0 else 0)
            for i in range(S)]
    cpu = None
    for cam in cams:
        cpu, _ = ol.oracle_update(cam, scene, W, H, accum=cpu)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update_n(cams)
        gpu = r.read_accum()
        # one more frame changes nothing: the sample cap is reached (compute_shader.wgsl:343, 352)
        r.update(scenes.camera(W, H, 30, S, random_seed=0.123, moved=0))
        capped = r.read_accum()
    assert (gpu[..., 3] == float(S)).all()
    assert_bit_identical(gpu, cpu, "default scene 128x72, depth 30, 500 spp")
    assert np.array_equal(capped.view(np.uint32), gpu.view(np.uint32))
    # converged: neighbouring sky pixels are smooth (a 1-spp frame is not) — the mean of 500 samples, not the last one
    assert float(np.abs(np.diff(gpu[0, :, :3], axis=0)).max()) < 0.02
