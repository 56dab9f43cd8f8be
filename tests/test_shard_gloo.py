"""world_size-2 gloo test (CPU) of the multi-GPU host logic: ray ranges partition the frame, each rank renders
its range, one gather assembles on rank 0 exactly the single-rank frame.  The per-range producer here is the CPU
oracle (test infrastructure); on GPUs bench.py runs the same shard.py code over NCCL with nerf_render_device."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerf_rs_b200 import shard

from .conftest import CAM_JSON, DATA

W, H, NC, NF = 16, 8, 8, 8


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    from oracle import pyoracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cam = O.camera_from_json(CAM_JSON)
        oc, of = O.Network(os.path.join(DATA, "coarse")), O.Network(os.path.join(DATA, "fine"))
        n_rays = W * H
        b, e = shard.ray_range(rank, world, n_rays)
        local = torch.from_numpy(O.render(oc, of, cam, W, H, NC, NF, ray_begin=b, ray_end=e, threads=1))
        frame = torch.zeros((n_rays, 3), dtype=torch.float32) if rank == 0 else None
        got = shard.gather_frame(local, frame, rank, world, dist)
        if rank == 0:
            np.save(out_path, got.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_ray_ranges_partition():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 7, 128, 640000):
            r = [shard.ray_range(k, world, n) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            assert sum(shard.ray_counts(world, n)) == n
            assert max(shard.ray_counts(world, n)) - min(shard.ray_counts(world, n)) <= 1
    assert shard.ray_range(7, 8, 640000) == (560000, 640000)
    with pytest.raises(ValueError):
        shard.ray_range(2, 2, 10)


@pytest.mark.parametrize("world", [2, 3])
def test_gather_equals_single_rank(tmp_path, onets, ocam, world):
    from oracle import pyoracle as O
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    want = O.render(onets[0], onets[1], ocam, W, H, NC, NF)
    assert (np.load(out) == want).all()  # bit for bit: a ray range renders exactly the rows of the full frame


def test_striped_views_partition_the_frame():
    """Every pixel belongs to exactly one part, in ascending order inside a part, for ragged sizes too."""
    import numpy as np
    from nerf_rs_b200 import shard
    for n, stripe, parts in ((640000, 800, 8), (640000, 800, 3), (1000, 64, 4), (10, 64, 4), (7, 1, 3), (129, 128, 2)):
        views = [shard.view_pixels(p, parts, stripe, n) for p in range(parts)]
        for v in views:
            assert (np.diff(v) > 0).all()
        allpix = np.sort(np.concatenate(views))
        assert (allpix == np.arange(n)).all()
        sizes = [len(v) for v in views]
        assert max(sizes) - min(sizes) <= stripe
    import pytest
    with pytest.raises(ValueError):
        shard.view_pixels(2, 2, 8, 100)
