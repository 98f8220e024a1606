"""Frame-parallel host logic under a real world_size-2 process group (gloo, CPU)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, out):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import pcc_b200
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = pcc_b200.frames.frames_of_rank(n_frames, rank, world)
        # pretend each frame t has 1000 + t points and takes (rank + 1) * 0.5 s in total on this rank
        pts = sum(1000 + t for t in mine)
        (t_dev, t_e2e), total = pcc_b200.frames.reduce_throughput([0.5 * (rank + 1), 0.25 * (world - rank)], pts, "cpu")
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            torch.save({"frames": gathered, "t_dev": t_dev, "t_e2e": t_e2e, "total": total}, out)
    finally:
        dist.destroy_process_group()


def test_frame_sharding_and_reduction_world2(tmp_path):
    world, n_frames = 2, 7
    out = str(tmp_path / "r.pt")
    mp.spawn(_worker, args=(world, _free_port(), n_frames, out), nprocs=world, join=True)
    r = torch.load(out)
    assert sorted(sum(r["frames"], [])) == list(range(n_frames))          # every frame exactly once
    assert r["frames"][0] == [0, 2, 4, 6] and r["frames"][1] == [1, 3, 5]
    assert r["t_dev"] == 1.0 and r["t_e2e"] == 0.5                          # max over ranks
    assert r["total"] == sum(1000 + t for t in range(n_frames))             # sum over ranks


def test_single_process_is_identity():
    import pcc_b200
    assert pcc_b200.frames.frames_of_rank(5, 0, 1) == [0, 1, 2, 3, 4]
    assert pcc_b200.frames.reduce_throughput([1.5], 10) == ([1.5], 10)
