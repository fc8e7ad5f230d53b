"""Host-side multi-GPU logic on CPU: world_size-2 gloo processes run the band / camera partition and the framebuffer
gather of gaussian-ray-tracing_b200/partition.py (the same code path bench.py uses over NCCL)."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pattern(H, W):
    y, x = np.mgrid[0:H, 0:W]
    return np.stack([(x * 3 + y) % 251, (x + y * 7) % 241, (x * y) % 239], axis=-1).astype(np.uint8)


def _worker(rank, world, port, H, W, band, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    part = importlib.import_module("gaussian-ray-tracing_b200.partition")
    full = _pattern(H, W)
    frame = np.zeros_like(full)
    for y0, rows in part.bands_for_rank(H, rank, world, band):      # "trace" only my bands
        frame[y0:y0 + rows] = full[y0:y0 + rows]
    out = part.gather_bands(torch.from_numpy(frame), rank, world, band)
    ok_bands = (rank != 0) or np.array_equal(out.numpy(), full)
    cams = part.cameras_for_rank(5, rank, world)
    mine = torch.full((4, 6, 3), 10 * rank + 1, dtype=torch.uint8)
    frames = part.gather_frames(mine, rank, world)
    ok_frames = (rank != 0) or all(int(frames[r][0, 0, 0]) == 10 * r + 1 for r in range(world))
    q.put((rank, bool(ok_bands), bool(ok_frames), cams, out is None))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("H,W,band", [(64, 48, 8), (70, 33, 8), (17, 5, 4)])
def test_band_partition_and_gather_world2(H, W, band):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + H) % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, H, W, band, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] and res[0][2] and not res[0][4]       # rank 0 assembled the exact frame
    assert res[1][4]                                       # rank 1 gets nothing back
    assert res[0][3] == [0, 2, 4] and res[1][3] == [1, 3]  # cameras k -> rank k mod G


def test_band_partition_covers_every_row_once():
    part = importlib.import_module("gaussian-ray-tracing_b200.partition")
    for H in (1, 7, 8, 9, 1080, 2160):
        for world in (1, 2, 4, 8):
            seen = np.zeros(H, int)
            for r in range(world):
                for y0, rows in part.bands_for_rank(H, r, world):
                    seen[y0:y0 + rows] += 1
                    assert np.all(part.band_owner_map(H, world)[y0:y0 + rows] == r)
            assert np.all(seen == 1)
