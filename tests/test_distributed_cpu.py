"""world_size-2 gloo tests of the N > 1 host logic (run on CPU): row-block and frame sharding plans
and the gather that reassembles them.  The oracle stands in for the kernel here — what is under test
is the partitioning/gather plumbing, which is identical on the GPU box (NCCL instead of gloo)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 48, 30   # 30 rows: 8 warp-tile rows, so the two blocks are 16 + 14 rows


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_lib as ol
    import scenes
    from gpu_ray_tracing_b200 import sharding
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=5, spp=500)
    # rows plan
    b, c = sharding.rows_for_rank(H, rank, world)
    block, _ = ol.oracle_update(cam, scene, W, H, row_begin=b, row_count=c)
    full = sharding.gather_row_blocks(dist, torch.from_numpy(block), H, rank, world)
    # bands plan: 8-row bands dealt round-robin; the oracle renders each band of this rank as a row window
    mine_rows = sharding.band_rows_for_rank(H, 8, rank, world)
    parts = [ol.oracle_update(cam, scene, W, H, row_begin=int(r0), row_count=int(min(8, H - r0)))[0] for r0 in mine_rows[::8]]
    banded = sharding.gather_bands(dist, torch.from_numpy(np.concatenate(parts)), H, 8, rank, world)
    # frames plan: 5 frames, frame k = random_seed k/8
    mine = sharding.frames_for_rank(5, rank, world)
    frames = {k: ol.oracle_update(scenes.camera(W, H, max_depth=2, spp=1, random_seed=k / 8.0), scene, W, H)[0] for k in mine}
    gathered = [None] * world
    dist.all_gather_object(gathered, {k: v.tobytes() for k, v in frames.items()})
    if rank == 0:
        np.save(os.path.join(out_dir, "rows.npy"), full.numpy())
        np.save(os.path.join(out_dir, "bands.npy"), banded.numpy())
        merged = {}
        for g in gathered:
            merged.update(g)
        np.save(os.path.join(out_dir, "frames.npy"), np.stack([np.frombuffer(merged[k], np.float32).reshape(H, W, 4) for k in range(5)]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharding_reassembles_the_single_process_result(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    import scenes
    ol.oracle()          # build before forking
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    scene, cam = scenes.default_scene(), scenes.camera(W, H, max_depth=5, spp=500)
    want, _ = ol.oracle_update(cam, scene, W, H)
    got = np.load(os.path.join(tmp_path, "rows.npy"))
    assert got.shape == want.shape and np.array_equal(got.view(np.uint32), want.view(np.uint32))
    got = np.load(os.path.join(tmp_path, "bands.npy"))
    assert got.shape == want.shape and np.array_equal(got.view(np.uint32), want.view(np.uint32))
    frames = np.load(os.path.join(tmp_path, "frames.npy"))
    for k in range(5):
        ref, _ = ol.oracle_update(scenes.camera(W, H, max_depth=2, spp=1, random_seed=k / 8.0), scene, W, H)
        assert np.array_equal(frames[k].view(np.uint32), ref.view(np.uint32)), k


def test_plans_cover_everything_once():
    sys.path.insert(0, ROOT)
    from gpu_ray_tracing_b200 import sharding
    for world in (1, 2, 4, 8):
        seen = sorted(k for r in range(world) for k in sharding.frames_for_rank(256, r, world))
        assert seen == list(range(256))
        rows = [sharding.rows_for_rank(4320, r, world) for r in range(world)]
        assert rows[0][0] == 0 and sum(c for _, c in rows) == 4320
        assert all(rows[i][0] + rows[i][1] == rows[i + 1][0] for i in range(world - 1))
        banded = np.concatenate([sharding.band_rows_for_rank(4320, 16, r, world) for r in range(world)])
        assert sorted(banded.tolist()) == list(range(4320))
