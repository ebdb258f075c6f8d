"""N > 1 on real GPUs: every rank renders its row tiles, ONE NCCL allgather, device de-interleave;
the assembled frame must be byte-identical to the single-GPU frame.
Needs >= 2 CUDA devices (skipped on the 1-GPU test box; run with `gpurun --gpus 2`)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    from simple_raytracer_b200 import scenegen
    from simple_raytracer_b200.multigpu import DistributedFrame
    from simple_raytracer_b200.renderer import DeviceScene

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        stream = torch.cuda.Stream()
        torch.cuda.set_stream(stream)
        for name, scene in (("soup", scenegen.stress_scene(20_000, 100, 320, 203, tex_size=64)),
                            ("coverage", scenegen.coverage_scene(160, 101))):
            with DeviceScene(scene) as ds:
                ds.render(rank=rank, world_size=world, tile_rows=10)          # sizes the queues
                df = DistributedFrame(ds, 10, rank, world)
                df.step(stream.cuda_stream)
                df.step(stream.cuda_stream)                                   # twice: no state leaks between frames
                torch.cuda.synchronize()
                np.save(os.path.join(out_dir, f"{name}_{rank}.npy"), df.frame_host())
                if rank == 0:
                    ds.render(tile_rows=10)
                    rgb, flags = ds.download()
                    np.save(os.path.join(out_dir, f"{name}_single.npy"), np.concatenate([rgb, flags[..., None]], axis=2))
    finally:
        dist.destroy_process_group()


def test_nccl_allgather_frame_equals_single_gpu_frame(tmp_path):
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 CUDA devices")
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for name in ("soup", "coverage"):
        want = np.load(tmp_path / f"{name}_single.npy")
        for rank in range(world):
            got = np.load(tmp_path / f"{name}_{rank}.npy")
            assert np.array_equal(got, want), f"{name}: rank {rank} assembled a different frame"
