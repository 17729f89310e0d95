"""Host-side logic of the multi-GPU path on CPU: world_size-2 gloo processes run the same GatherFrame code
bench.py uses under NCCL (partitioning into interleaved tiles, padding, gather, de-interleave)."""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, width, height, tile_rows, out_path):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mg = importlib.import_module("ray-tracer-challenge_b200.multi_gpu")
        g = mg.GatherFrame(width, height, tile_rows, rank, world, torch.device("cpu"))
        ys = mg.partition_rows(height, tile_rows, rank, world)
        # stand-in for the render: row y of the frame holds y + 1 in R, rank in G, x in B
        for k, y in enumerate(ys):
            g.local[k, :, 0] = y + 1
            g.local[k, :, 1] = rank
            g.local[k, :, 2] = torch.arange(width, dtype=torch.float64)
        frame = g.assemble()
        if rank == 0:
            torch.save(frame, out_path)
        else:
            assert frame is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("height,tile_rows", [(37, 8), (16, 8), (5, 8), (64, 1)])
def test_gather_frame_world2_gloo(tmp_path, height, tile_rows):
    world, width = 2, 11
    out = str(tmp_path / "frame.pt")
    mp.spawn(_worker, args=(world, _free_port(), width, height, tile_rows, out), nprocs=world, join=True)
    frame = torch.load(out)
    assert frame.shape == (height, width, 3)
    for y in range(height):
        assert torch.all(frame[y, :, 0] == y + 1)
        assert torch.all(frame[y, :, 1] == (y // tile_rows) % world)
        assert torch.all(frame[y, :, 2] == torch.arange(width, dtype=torch.float64))


def test_partition_rows_matches_c_abi():
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    mg = importlib.import_module("ray-tracer-challenge_b200.multi_gpu")
    cam = pkg.rtc_camera()
    for vsize in (1, 7, 8, 9, 64, 1080):
        cam.vsize, cam.hsize = vsize, 4
        c = pkg.Camera(cam)
        for n in (1, 2, 3, 8):
            seen = []
            for r in range(n):
                ys = mg.partition_rows(vsize, 8, r, n)
                assert len(ys) == c.partition_rows((8, r, n))
                seen += ys
            assert sorted(seen) == list(range(vsize))  # a partition: every row exactly once


def _shared_worker(rank, world, port, width, height, tile_rows, out_path):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mg = importlib.import_module("ray-tracer-challenge_b200.multi_gpu")
        f = mg.SharedHostFrame(width, height, rank, world, register=False)  # no GPU here: the mapping logic only
        for y in mg.partition_rows(height, tile_rows, rank, world):  # stand-in for the render of this rank's rows
            f.array[y, :, 0] = y + 1
            f.array[y, :, 1] = rank
        f.barrier()  # the canvas's own shared-memory barrier (what bench.py's N-GPU end-to-end leg ends a frame with)
        if rank == 0:
            torch.save(torch.from_numpy(f.array.copy()), out_path)
        for _ in range(3):
            f.barrier()
        dist.barrier()
        f.close()
    finally:
        dist.destroy_process_group()


def test_shared_host_frame_world2_gloo(tmp_path):
    """Two processes write their interleaved rows into ONE host canvas (the N-GPU end-to-end path of bench.py)."""
    world, width, height, tile_rows = 2, 9, 21, 4
    out = str(tmp_path / "frame.pt")
    mp.spawn(_shared_worker, args=(world, _free_port(), width, height, tile_rows, out), nprocs=world, join=True)
    frame = torch.load(out)
    for y in range(height):
        assert torch.all(frame[y, :, 0] == y + 1)
        assert torch.all(frame[y, :, 1] == (y // tile_rows) % world)


def _bcast_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pkg = importlib.import_module("ray-tracer-challenge_b200")
        mg = importlib.import_module("ray-tracer-challenge_b200.multi_gpu")
        from tests import scenes
        blob = cam = None
        if rank == 0:  # only rank 0 parses the scene
            blob, cam = pkg.load_blob(scenes.SMOKE_SCENE, "", 96, 64)
        blob, cam = mg.broadcast_scene(blob, cam, rank, torch.device("cpu"))
        info = pkg.inspect_blob(blob)  # every rank can build the device scene from what it received
        with open(os.path.join(out_dir, "r%d.txt" % rank), "w") as f:
            import hashlib
            f.write("%s %d %d %r %d" % (hashlib.sha256(blob).hexdigest(), cam.hsize, cam.vsize, cam.pixel_size, info.n_leaves))
    finally:
        dist.destroy_process_group()


def test_scene_broadcast_world2_gloo(tmp_path):
    """The N > 1 path parses the scene on rank 0 only and broadcasts blob + camera (NCCL on GPUs, gloo here)."""
    mp.spawn(_bcast_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    a, b = (open(os.path.join(str(tmp_path), "r%d.txt" % r)).read() for r in range(2))
    assert a == b and a.split()[1:3] == ["96", "64"]
