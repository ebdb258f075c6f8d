"""The N > 1 frame assembly on CPU: world_size-2 (and 3) `gloo` process groups exchange
tile buffers exactly as the GPU ranks do over NCCL (simple_raytracer_b200/multigpu.py).
The pixels come from the CPU oracle here (no GPU in this suite); what is under test is
the sharding arithmetic of the C library (rt_tile_layout), the single allgather and the
de-interleave — the host-side logic of SURVEY 8e."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from simple_raytracer_b200 import multigpu, scenegen
from tests import oracle_lib as O


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _frame_rgba(height, width):
    """A frame with real content: the oracle's render of the coverage scene, RGBA8 with flags in A."""
    scene = scenegen.coverage_scene(width, height)
    raw, ids, _ = O.oracle_render(scene)
    rgb, flags = O.raw_to_rgb8(raw, ids)
    return np.concatenate([rgb.reshape(height, width, 3), flags.reshape(height, width, 1)], axis=2)


def _worker(rank, world, port, height, width, tile_rows, frame, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lay = multigpu.tile_layout(height, width, world, tile_rows)
        # this rank "renders" only the rows it owns
        send = torch.from_numpy(multigpu.shard_host(frame, lay, rank).reshape(-1).copy())
        assert send.numel() == lay.bytes_per_rank
        gathered = multigpu.allgather_tiles(send)
        full = multigpu.deinterleave_host(gathered.numpy(), lay)
        np.save(os.path.join(out_dir, f"frame_{rank}.npy"), full)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,height,width,tile_rows", [(2, 61, 40, 8), (2, 64, 32, 8), (3, 50, 24, 5)])
def test_allgather_of_row_tiles_reassembles_the_frame(tmp_path, world, height, width, tile_rows):
    frame = _frame_rgba(height, width)
    port = _free_port()
    mp.spawn(_worker, args=(world, port, height, width, tile_rows, frame, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        got = np.load(tmp_path / f"frame_{rank}.npy")
        assert np.array_equal(got, frame), f"rank {rank} assembled a different frame"


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("height,tile_rows", [(2160, 10), (4320, 10), (1080, 5), (101, 8), (7, 8)])
def test_tile_layout_partitions_the_rows(world, height, tile_rows):
    lay = multigpu.tile_layout(height, 16, world, tile_rows)
    assert lay.n_tiles == -(-height // tile_rows)
    assert lay.tiles_per_rank == -(-lay.n_tiles // world)
    assert lay.rows_per_rank == lay.tiles_per_rank * tile_rows and lay.pixels_per_rank == lay.rows_per_rank * 16
    seen = np.concatenate([lay.rows_of_rank(r) for r in range(world)])
    seen = seen[seen >= 0]
    assert sorted(seen.tolist()) == list(range(height))      # every row exactly once
    # interleaving: consecutive tiles go to consecutive ranks
    for r in range(world):
        rows = lay.rows_of_rank(r)
        if rows[0] >= 0:
            assert rows[0] == r * tile_rows
