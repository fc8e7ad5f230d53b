"""Multi-GPU partition of the hot path: rays are independent and the scene is read-only, so every rank holds
the whole scene + BVH and the work is split with NO data-path collective; the only exchange is the framebuffer gather.

Two partitions, as north_star names them:
  * camera batches (BASELINE config 5): camera k -> rank k mod G;
  * image tiles of one camera: horizontal bands of `band` rows, interleaved over the ranks for load balance.
torch.distributed is plumbing only (NCCL on the GPU box, gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Tuple


def cameras_for_rank(n_cameras: int, rank: int, world: int) -> List[int]:
    return list(range(rank, n_cameras, world))


def bands_for_rank(height: int, rank: int, world: int, band: int = 8) -> List[Tuple[int, int]]:
    """[(y0, rows), ...] of this rank: band b (rows [b*band, (b+1)*band)) belongs to rank b mod world."""
    out = []
    nb = (height + band - 1) // band
    for b in range(rank, nb, world):
        y0 = b * band
        out.append((y0, min(band, height - y0)))
    return out


def band_owner_map(height: int, world: int, band: int = 8):
    """For every image row the rank that traces it (used by the gather to put rows back)."""
    import numpy as np
    rows = np.arange(height)
    return (rows // band) % world


def gather_bands(frame, rank: int, world: int, band: int = 8, dst: int = 0):
    """Every rank passes its (H, W, 3) uint8 frame in which only its own bands are valid; rank `dst` returns the
    assembled frame, the others None.  One gather of the packed band rows (H/world rows per rank)."""
    import torch
    import torch.distributed as dist
    H = frame.shape[0]
    owner = torch.as_tensor(band_owner_map(H, world, band), device=frame.device)
    mine = frame[owner == rank].contiguous()
    # ranks may own different row counts when H is not a multiple of band*world: pad to the maximum
    counts = [int((owner == r).sum()) for r in range(world)]
    pad = max(counts)
    buf = torch.zeros((pad,) + tuple(frame.shape[1:]), dtype=frame.dtype, device=frame.device)
    buf[: mine.shape[0]] = mine
    if rank == dst:
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.gather(buf, parts, dst=dst)
        out = torch.empty_like(frame)
        for r in range(world):
            out[owner == r] = parts[r][: counts[r]]
        return out
    dist.gather(buf, None, dst=dst)
    return None


def gather_frames(frame, rank: int, world: int, dst: int = 0):
    """Camera-batch partition: every rank contributes one whole frame; rank `dst` gets the list."""
    import torch
    import torch.distributed as dist
    if rank == dst:
        parts = [torch.empty_like(frame) for _ in range(world)]
        dist.gather(frame, parts, dst=dst)
        return parts
    dist.gather(frame, None, dst=dst)
    return None
