"""GPU parity: the CUDA kernels, called through the C ABI, against the CPU oracle on the same
bytes.  The bar is bit-identical rgba32f accumulators (stronger than the <= 1/255 after RGBA8
quantisation the port target states); the RGBA8 mismatch count is reported either way."""
import numpy as np
import pytest

import gpu_ray_tracing_b200 as rt
import oracle_lib as ol
import scenes

pytestmark = pytest.mark.gpu


def gpu_update(scene, cams, width, height, variant=0, accum=None, row_begin=0, row_count=None):
    with rt.Renderer(width, height, row_begin=row_begin, row_count=row_count) as r:
        r.set_kernel_variant(variant)
        r.set_scene(scene)
        if accum is None:
            r.init()
        else:
            r.write_accum(accum)
        for cam in cams:
            r.update(cam)
        return r.read_accum(), r.rays_traced(), r.exact_tests()


def assert_parity(gpu, cpu, what):
    rep = ol.compare_images(cpu, gpu)
    print(what, rep)
    assert rep["pixels_over_1_255"] == 0, (what, rep)
    assert rep["bit_identical"], (what, rep)


@pytest.mark.parametrize("variant", [1, 2, 3, 4])
@pytest.mark.parametrize("depth", [1, 30])
def test_default_scene_720p(variant, depth):
    """BASELINE config 1: repo default scene at 1280x720, first frame."""
    W, H = 1280, 720
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=depth, spp=500)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H, variant)
    assert rays_gpu == rays_cpu
    assert_parity(gpu, cpu, f"default 720p depth {depth} variant {variant}")


@pytest.mark.parametrize("variant", [2, 3, 4])
def test_default_scene_4k_depth1(variant):
    """BASELINE config 2: default scene at 3840x2160 (layout + tolerance check)."""
    W, H = 3840, 2160
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=1, spp=1)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H, variant)
    assert rays_gpu == rays_cpu == W * H
    assert_parity(gpu, cpu, f"default 4K variant {variant}")


@pytest.mark.parametrize("defocus", [None, 0.0])
@pytest.mark.parametrize("variant", [1, 2, 3, 4])
def test_random_1024_spheres(variant, defocus):
    """BASELINE config 3 scene at a size the oracle finishes in seconds (default camera, and a pinhole)."""
    W, H = 960, 540
    scene, cam = scenes.random_scene(1024), scenes.camera(W, H, max_depth=1, spp=1, defocus_angle=defocus)
    cpu, _ = ol.oracle_update(cam, scene, W, H)
    gpu, rays, exact = gpu_update(scene, [cam], W, H, variant)
    assert rays == W * H
    assert_parity(gpu, cpu, f"random1024 variant {variant}")


def test_random_1024_depth8_progressive():
    """Three progressive frames with a re-rolled random_seed and the reset flag cleared."""
    W, H = 640, 360
    scene = scenes.random_scene(1024)
    cams = [scenes.camera(W, H, max_depth=8, spp=500, random_seed=s, moved=m)
            for s, m in ((0.25, 1), (0.625, 0), (0.8125, 0))]
    cpu = np.zeros((H, W, 4), dtype=np.float32)
    rays_cpu = 0
    for cam in cams:
        cpu, r = ol.oracle_update(cam, scene, W, H, accum=cpu)
        rays_cpu += r
    gpu, rays_gpu, _ = gpu_update(scene, cams, W, H)
    assert rays_gpu == rays_cpu
    assert (gpu[..., 3] == 3.0).all()
    assert_parity(gpu, cpu, "random1024 depth8 x3 frames")


def test_ragged_sizes_and_windows():
    """Sizes that are not multiples of the 8x8 workgroup / 32x4 warp tile, and a row window."""
    scene = scenes.default_scene()
    for (W, H) in ((33, 7), (1, 1), (250, 131)):
        cam = scenes.camera(W, H, max_depth=4, spp=2)
        cpu, _ = ol.oracle_update(cam, scene, W, H)
        for variant in (1, 2, 3, 4):
            gpu, _, _ = gpu_update(scene, [cam], W, H, variant)
            assert_parity(gpu, cpu, f"{W}x{H} variant {variant}")
    W, H = 250, 131
    cam = scenes.camera(W, H, max_depth=4, spp=2)
    cpu, _ = ol.oracle_update(cam, scene, W, H)
    gpu, _, _ = gpu_update(scene, [cam], W, H, row_begin=37, row_count=50)
    assert_parity(gpu, cpu[37:87], "row window 37..87")


def test_sample_cap_and_empty_scene():
    """n >= samples_per_pixel leaves the texel alone; zero spheres renders pure sky; max_depth 0 too."""
    W, H = 64, 32
    scene = scenes.default_scene()
    cam = scenes.camera(W, H, max_depth=5, spp=2, moved=0)
    start = np.zeros((H, W, 4), dtype=np.float32)
    start[..., :3] = 0.25
    start[..., 3] = 2.0
    gpu, rays, _ = gpu_update(scene, [cam], W, H, accum=start.copy())
    assert rays == 0 and np.array_equal(gpu, start)
    empty = rt.SphereCollection(np.zeros(4, dtype=rt.SPHERE_DTYPE), 0)
    cam = scenes.camera(W, H, max_depth=5, spp=1)
    cpu, _ = ol.oracle_update(cam, empty, W, H)
    gpu, _, _ = gpu_update(empty, [cam], W, H)
    assert_parity(gpu, cpu, "empty scene")
    cam0 = scenes.camera(W, H, max_depth=0, spp=1)
    cpu, rays_cpu = ol.oracle_update(cam0, scene, W, H)
    gpu, rays_gpu, _ = gpu_update(scene, [cam0], W, H)
    assert rays_cpu == rays_gpu == 0
    assert_parity(gpu, cpu, "max_depth 0")


def test_large_scene_streamed():
    """16,384 spheres (four shared-memory chunks per traversal) — the config-5 path at small size."""
    W, H = 256, 144
    scene = scenes.grid_scene(16384)
    cam = rt.orbit_camera(3, 256, W, H)
    cpu, _ = ol.oracle_update(cam, scene, W, H)
    gpu, rays, _ = gpu_update(scene, [cam], W, H)
    assert rays == W * H
    assert_parity(gpu, cpu, "grid 16384")
    gpu3, _, _ = gpu_update(scene, [cam], W, H, variant=3)
    assert_parity(gpu3, cpu, "grid 16384 general streamed")
    cam.max_depth = 6.0
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H)
    assert rays_cpu == rays_gpu
    assert_parity(gpu, cpu, "grid 16384 depth 6 (tree walked in global memory)")
    gpu5, rays_gpu5, _ = gpu_update(scene, [cam], W, H, variant=5)
    assert rays_cpu == rays_gpu5
    assert_parity(gpu5, cpu, "grid 16384 depth 6 (streamed bounce kernel)")
    # 3000 spheres: three chunks with a ragged last one through the streamed primary pass, defocus camera
    scene = scenes.random_scene(3000)
    cam = scenes.camera(W, H, max_depth=3, spp=2)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H, variant=4)
    assert rays_cpu == rays_gpu
    assert_parity(gpu, cpu, "random 3000 streamed primary")


def test_rgba8_and_headless_render():
    W, H = 320, 180
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=10, spp=4)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        accum, rgba = r.read_accum(), r.read_rgba8()
    assert np.array_equal(rgba, ol.quantize(accum))
    out = rt.render(scene, cam, W, H, samples=1)
    assert np.array_equal(out, accum)


def test_node_state_machine_on_device():
    """Loading -> Init -> Update(1) <-> Update(0): the Init dispatch zeroes, every later one samples."""
    W, H = 64, 64
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=3, spp=10, moved=0)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        node = rt.ComputeShaderNode(r)
        counts = []
        for _ in range(5):
            node.update()
            node.run(cam)
            counts.append(float(r.read_accum()[0, 0, 3]))
    assert counts == [0.0, 1.0, 2.0, 3.0, 4.0]


@pytest.mark.parametrize("n_shards", [2, 3])
def test_multi_renderer_row_shards(n_shards):
    """One image sharded by row blocks (config 4 shape): gathered output == the single-renderer image.  Uses as many
    distinct devices as the box has (a single GPU hosts all shards; the peer copies become local copies)."""
    W, H = 320, 182
    scene = scenes.random_scene(256)
    have = rt.device_count()
    devices = [i % have for i in range(n_shards)]
    for kw in (dict(max_depth=1, spp=1), dict(max_depth=5, spp=3)):
        cam = scenes.camera(W, H, **kw)
        single, rays_single, _ = gpu_update(scene, [cam], W, H)
        with rt.MultiRenderer(devices, W, H) as m:
            m.set_scene(scene)
            m.init()
            m.update(cam)
            accum = m.gather_accum()
            rgba = m.gather_rgba8()
            assert m.rays_traced() == rays_single
        assert np.array_equal(accum.view(np.uint32), single.view(np.uint32))
        assert np.array_equal(rgba, ol.quantize(single))


def test_async_readback_matches_blocking_read():
    """rt_read_rgba8_async: three frames in flight through two device slots, each equal to the blocking read."""
    import torch
    W, H = 256, 144
    scene = scenes.default_scene()
    cams = [scenes.camera(W, H, max_depth=3, spp=4, random_seed=s, moved=m) for s, m in ((0.125, 1), (0.5, 0), (0.75, 0), (0.875, 1))]
    want = []
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        for cam in cams:
            r.update(cam)
            want.append(r.read_rgba8().copy())
    bufs = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in cams]
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        for cam, buf in zip(cams, bufs):
            r.update(cam)
            r.read_rgba8_async_into(buf.data_ptr())
        r.read_wait()
    for w, buf in zip(want, bufs):
        assert np.array_equal(w, buf.numpy())


def test_wavefront_counts_and_depth_fallback():
    """Deep frames: the per-bounce queue lengths add up to the rays the oracle traces; max_depth beyond the
    wavefront's 64-bounce bookkeeping takes the in-kernel bounce loop and still matches."""
    W, H = 200, 120
    scene = scenes.default_scene()
    cam = scenes.camera(W, H, max_depth=12, spp=500)
    cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.update(cam)
        gpu, rays_gpu, counts = r.read_accum(), r.rays_traced(), r.wavefront_counts()
    assert_parity(gpu, cpu, "wavefront depth 12")
    assert rays_gpu == rays_cpu == W * H + sum(counts[1:12]) and counts[12] == 0 and counts[1] > 0
    for depth in (64, 65):
        cam = scenes.camera(W, H, max_depth=depth, spp=500, defocus_angle=0.0)
        cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
        gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H)
        assert rays_gpu == rays_cpu
        assert_parity(gpu, cpu, f"depth {depth}")


def test_scripted_flythrough_frames():
    """the keyboard controller scripted (yaw, dolly, an idle frame that accumulates instead of
    resetting); every frame's accumulator equals the oracle's fed the same 176-byte uniforms."""
    W, H, dt = 320, 200, 1.0 / 30.0
    K = rt.Key
    keys = [K.ARROW_LEFT, K.ARROW_LEFT | K.W, 0, 0, K.DIGIT1, K.S | K.D, 0]
    settings = rt.CameraSettings.default()
    settings.max_depth = 6
    cams = rt.camera_script(settings, keys, dt, W, H, seeds=[0.0625 * (f + 1) for f in range(len(keys))])
    assert [c.camera_has_moved for c in cams] == [1.0, 1.0, 0.0, 0.0, 1.0, 1.0, 0.0]
    scene = scenes.default_scene()
    cpu = None
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        for f, cam in enumerate(cams):
            cpu, _ = ol.oracle_update(cam, scene, W, H, accum=cpu)
            r.update(cam)
            gpu = r.read_accum()
            assert_parity(gpu, cpu, f"fly-through frame {f}")
            assert gpu[..., 3].max() == (1.0 if cam.camera_has_moved else gpu[..., 3].max())
        assert cpu[..., 3].max() == 2.0     # frames 5, 6: one moved frame then one idle frame


@pytest.mark.parametrize("variant", [0, 2, 3])
def test_degenerate_spheres_on_device(variant):
    """NaN / infinite / out-of-window spheres next to ordinary ones: every kernel family must follow the reference
    loop, including its order-dependent behaviour once a root is NaN (tree traversal: sequential retrace)."""
    scene = scenes.degenerate_scene()
    W, H = 64, 36
    for depth in (1, 5):
        cam = scenes.camera(W, H, max_depth=depth, spp=500, look_from=(9.0, 1.5, 2.0), fov=50.0)
        cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
        gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H, variant)
        assert rays_gpu == rays_cpu
        assert ol.nan_aware_equal(cpu, gpu), (variant, depth)


def test_tree_sizes_and_streamed_fallback():
    """Bounce rays through trees of different depths (8 ... 3,000 spheres: image in shared memory; 6,000 and 40,000:
    past that budget, walked in global memory with 32-bit candidate codes) and through the streamed-record kernel
    (variant 5); the counters of every path agree with the oracle's ray count."""
    W, H = 192, 108
    for n in (8, 9, 64, 65, 500, 3000, 6000, 40000):
        scene = scenes.random_scene(n) if n < 10000 else scenes.grid_scene(n)
        cam = scenes.camera(W, H, max_depth=5, spp=500)
        cpu, rays_cpu = ol.oracle_update(cam, scene, W, H)
        gpu, rays_gpu, _ = gpu_update(scene, [cam], W, H)
        assert rays_gpu == rays_cpu, n
        assert_parity(gpu, cpu, f"random {n} depth 5")
        if n in (500, 6000):
            gpu5, rays_gpu5, _ = gpu_update(scene, [cam], W, H, variant=5)
            assert rays_gpu5 == rays_cpu, n
            assert_parity(gpu5, cpu, f"random {n} depth 5, streamed bounce kernel")


def test_bounce_path_selection():
    """Which traversal the bounce rays get (rt_tree_info): shared-memory image for small scenes, the tree walked in
    global memory past that budget (config 5's 65,536 spheres included), streamed records only on request."""
    W, H = 64, 36
    for scene, variant, want in ((scenes.random_scene(1024), 0, 1), (scenes.random_scene(6000), 0, 2),
                                 (scenes.grid_scene(65536), 0, 2), (scenes.random_scene(6000), 5, 3)):
        with rt.Renderer(W, H) as r:
            r.set_kernel_variant(variant)
            r.set_scene(scene)
            r.init()
            assert r.tree_info() == (0, 0, 0)
            r.update(scenes.camera(W, H, max_depth=3, spp=1))
            nodes, depth, path = r.tree_info()
            assert nodes > scene.count // 8 and depth >= 2 and path == want, (scene.count, variant, nodes, depth, path)


def test_fused_rgba8_output():
    """rt_set_rgba8_output: depth-1 updates write the RGBA8 frame themselves; asynchronous reads of progressive frames,
    of a deep frame in between (quantise kernel again) and after rt_init equal the blocking read every time, and the
    fused frames cost no extra kernel."""
    import torch
    W, H = 250, 131
    scene = scenes.random_scene(300)
    host = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
    with rt.Renderer(W, H) as r:
        r.set_scene(scene)
        r.init()
        r.set_rgba8_output(True)
        seq = [(1, 1), (1, 0), (1, 0), (4, 1), (1, 1), (1, 0)]     # (max_depth, camera_has_moved)
        for i, (depth, moved) in enumerate(seq):
            cam = scenes.camera(W, H, max_depth=depth, spp=8, random_seed=(i + 1) / 16.0, moved=moved)
            before = r.kernel_launches()
            r.update(cam)
            r.read_rgba8_async_into(host[i & 1].data_ptr())
            used = r.kernel_launches() - before
            r.read_wait()
            want = r.read_rgba8()
            assert np.array_equal(host[i & 1].numpy(), want), (i, depth)
            if depth == 1:
                assert used == 2, used      # per-frame records + primary pass, no quantise kernel
        r.init()
        r.read_rgba8_async_into(host[0].data_ptr())
        r.read_wait()
        assert not host[0].numpy()[..., :3].any() and (host[0].numpy()[..., 3] == 255).all()
        # against the oracle's quantisation of its own image
        cam = scenes.camera(W, H, max_depth=1, spp=1)
        r.update(cam)
        r.read_rgba8_async_into(host[1].data_ptr())
        r.read_wait()
        cpu, _ = ol.oracle_update(cam, scene, W, H)
        assert np.array_equal(host[1].numpy(), ol.quantize(cpu))


def test_async_read_into_device_memory():
    """rt_read_rgba8_async with a DEVICE destination (what a peer process maps through CUDA IPC in bench.py --gpus N):
    the copy lands in a buffer from rt_device_alloc and equals the blocking host read."""
    import torch
    W, H = 320, 180
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=3, spp=4)
    nbytes = W * H * 4
    dst = rt.device_alloc(0, 2 * nbytes)
    try:
        handle = rt.ipc_export(0, dst)
        assert len(handle) == 64 and any(handle)
        with rt.Renderer(W, H) as r:
            r.set_scene(scene)
            r.init()
            r.update(cam)
            r.read_rgba8_async_into(dst + nbytes)          # second slot of the buffer
            want = r.read_rgba8()
            r.read_wait()
            # bring the device copy back through torch (same device, raw pointer wrapped by a pinned host read)
            host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
            from cuda.bindings import runtime as cudart
            err, = cudart.cudaMemcpy(host.data_ptr(), dst + nbytes, nbytes, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
            assert int(err) == 0
            assert np.array_equal(host.numpy().reshape(H, W, 4), want)
    finally:
        rt.device_free(0, dst)


def test_back_to_back_frames_keep_their_own_records():
    """Frames enqueued without a synchronisation between them: a depth-1 frame's per-frame records are prepared on a
    stream of their own WHILE the frame before it renders (double-buffered by frame parity), deep frames prepare theirs
    in line, and a scene upload sits in the middle.  Every frame ACCUMULATES (camera_has_moved = 0 after the first) from a
    pose of its own, so one frame traced with another frame's records would show in the running mean: the result must be
    the oracle's fed the same sequence, bit for bit."""
    W, H = 640, 360
    depths = [1, 1, 1, 5, 1, 1, 3, 1, 1, 1]
    cams = []
    for f, depth in enumerate(depths):
        cam = rt.orbit_camera(3 * f + 1, 64, W, H, radius=14.0, cam_height=2.5 + 0.2 * f, fov_deg=30.0)
        cam.max_depth = float(depth)
        cam.samples_per_pixel = 500.0
        cam.camera_has_moved = 1.0 if f == 0 else 0.0
        cam.random_seed = 0.03125 * (f + 1)
        cams.append(cam)
    scene_a, scene_b = scenes.default_scene(), scenes.random_scene(300)
    cpu = None
    for f, cam in enumerate(cams):
        cpu, _ = ol.oracle_update(cam, scene_a if f < 6 else scene_b, W, H, accum=cpu)
    with rt.Renderer(W, H) as r:
        r.set_scene(scene_a)
        r.init()
        for f, cam in enumerate(cams):
            if f == 6:
                r.set_scene(scene_b)
            r.update(cam)
        gpu = r.read_accum()
    assert_parity(gpu, cpu, "ten frames back to back, mixed depths, scene change")
    assert (gpu[..., 3] == float(len(cams))).all()
