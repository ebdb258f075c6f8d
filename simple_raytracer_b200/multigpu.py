"""Multi-GPU frame assembly: interleaved row tiles + one allgather.

One process per GPU.  Rank r of N renders the tiles t with t % N == r into a
rank-contiguous RGBA8 buffer `[local_tile][row][col]` (rt_render_async), the
buffers are exchanged with ONE `all_gather_into_tensor` (NCCL over NVLink on the
GPU box, gloo in the CPU tests) and put into scanline order (on the device:
rt_deinterleave_tiles; for a frame gathered into host memory: deinterleave_host).

The sharding arithmetic is the C library's (rt_tile_layout), so the Python side
and the kernels cannot disagree about which rank owns which row.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import abi


@dataclass(frozen=True)
class TileLayout:
    height: int
    width: int
    world_size: int
    tile_rows: int
    n_tiles: int          # tiles in the image
    tiles_per_rank: int   # padded: every rank holds the same count (allgather needs equal sizes)
    rows_per_rank: int
    pixels_per_rank: int

    @property
    def bytes_per_rank(self) -> int:
        return self.pixels_per_rank * 4       # RGBA8, A = flags

    def global_row(self, rank: int, local_row: int) -> int:
        """Image row of row `local_row` of `rank`'s tile buffer (may be >= height: padding)."""
        tile, in_tile = divmod(local_row, self.tile_rows)
        return (tile * self.world_size + rank) * self.tile_rows + in_tile

    def rows_of_rank(self, rank: int) -> np.ndarray:
        """Image rows rendered by `rank`, in the order they sit in its tile buffer; -1 = padding."""
        r = np.array([self.global_row(rank, k) for k in range(self.rows_per_rank)], dtype=np.int64)
        return np.where(r < self.height, r, -1)


def tile_layout(height: int, width: int, world_size: int, tile_rows: int) -> TileLayout:
    out = (C.c_int32 * 4)()
    abi.check(abi.load_rt().rt_tile_layout(height, width, world_size, tile_rows, out), "rt_tile_layout")
    return TileLayout(height, width, world_size, tile_rows, out[0], out[1], out[2], out[3])


def deinterleave_host(gathered: np.ndarray, layout: TileLayout) -> np.ndarray:
    """[rank][local_tile][row][col] RGBA8 (uint8, flat or shaped) -> [height][width][4]."""
    g = np.asarray(gathered, dtype=np.uint8).reshape(layout.world_size, layout.rows_per_rank, layout.width, 4)
    out = np.zeros((layout.height, layout.width, 4), dtype=np.uint8)
    for rank in range(layout.world_size):
        rows = layout.rows_of_rank(rank)
        ok = rows >= 0
        out[rows[ok]] = g[rank][ok]
    return out


def shard_host(frame_rgba: np.ndarray, layout: TileLayout, rank: int) -> np.ndarray:
    """The tile buffer rank `rank` would produce for a given full frame (padding rows zero)."""
    rows = layout.rows_of_rank(rank)
    buf = np.zeros((layout.rows_per_rank, layout.width, 4), dtype=np.uint8)
    ok = rows >= 0
    buf[ok] = frame_rgba[rows[ok]]
    return buf


def allgather_tiles(send, gathered=None, group=None):
    """One collective per frame.  `send`: this rank's tile buffer as a flat uint8 torch
    tensor (CUDA -> NCCL, CPU -> gloo).  Returns the [world * bytes_per_rank] tensor."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if gathered is None:
        gathered = torch.empty(send.numel() * world, dtype=torch.uint8, device=send.device)
    dist.all_gather_into_tensor(gathered, send, group=group)
    return gathered


class DistributedFrame:
    """Per-rank state of the multi-GPU frame loop on a CUDA device (used by bench.py)."""

    def __init__(self, device_scene, tile_rows: int, rank: int, world_size: int, max_depth: int = 4,
                 epsilon: float = 1.0e-3, accel: int = abi.RT_ACCEL_AUTO):
        import torch
        from .renderer import make_params
        self.ds = device_scene
        self.rank, self.world, self.tile_rows = rank, world_size, tile_rows
        sc = device_scene.scene
        self.layout = tile_layout(sc.height, sc.width, world_size, tile_rows)
        self.params = make_params(max_depth=max_depth, epsilon=epsilon, accel=accel, rank=rank,
                                  world_size=world_size, tile_rows=tile_rows)
        n = self.layout.bytes_per_rank
        assert n == device_scene.tile_buffer_bytes(world_size, tile_rows)
        self.send = torch.empty(n, dtype=torch.uint8, device="cuda")
        self.gathered = torch.empty(n * world_size, dtype=torch.uint8, device="cuda") if world_size > 1 else self.send
        self.frame = torch.empty(sc.height * sc.width * 4, dtype=torch.uint8, device="cuda")

    def step(self, stream_ptr: int) -> None:
        """Enqueue one frame on the CUDA stream `stream_ptr`: render this rank's tiles, allgather, de-interleave.
        The collective is issued with that stream current (torch orders NCCL work against the CURRENT stream),
        so the three steps are stream-ordered whatever stream the caller had current."""
        import torch
        self.ds.render_async(self.params, self.send.data_ptr(), stream_ptr)
        if self.world > 1:
            with torch.cuda.stream(torch.cuda.ExternalStream(stream_ptr)):
                allgather_tiles(self.send, self.gathered)
        self.ds.deinterleave(self.gathered.data_ptr(), self.frame.data_ptr(), self.world, self.tile_rows, stream_ptr)

    def frame_host(self) -> np.ndarray:
        sc = self.ds.scene
        return self.frame.cpu().numpy().reshape(sc.height, sc.width, 4)
