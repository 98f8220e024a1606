"""Slab partition host logic on CPU: the neighbour exchange / gather semantics of both communicators
(torch.distributed over gloo with world_size 2 and 3; virtual ranks as threads) and the boundary cut."""
import os
import socket
import threading

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_rows(rank):
    """Rank r owns 4 + r rows [r * 100 + i, ...] x 3 columns; sends its first (1 + r) rows left and
    its last 2 rows right."""
    n = 4 + rank
    F = (torch.arange(n, dtype=torch.float32)[:, None] + 100.0 * rank) * torch.ones(1, 3)
    return F, 1 + rank, 2


def _scenario(comm):
    rank, world = comm.rank, comm.world
    F, nl, nr = _rank_rows(rank)
    n = F.shape[0]
    hl = 2 if rank > 0 else 0                    # what the left neighbour sends right
    hr = (1 + rank + 1) if rank < world - 1 else 0  # what the right neighbour sends left
    local = torch.full((n + hl + hr, 3), -1.0)
    local[:n] = F
    comm.exchange(local[:nl], local[n - nr:n], local[n:n + hl], local[n + hl:n + hl + hr])
    allrows, counts = comm.all_gather_rows(F[:, 0].contiguous().long())
    total = comm.all_reduce_sum(n, "cpu")
    keys = torch.empty(0, dtype=torch.int64)
    empty, ecounts = comm.all_gather_rows(keys if rank else torch.tensor([7], dtype=torch.int64))
    return {"local": local, "all": allrows, "counts": counts, "total": total, "empty": empty, "ecounts": ecounts}


def _check(results, world):
    for rank, r in enumerate(results):
        F, _, _ = _rank_rows(rank)
        n = F.shape[0]
        assert torch.equal(r["local"][:n], F)
        pos = n
        if rank > 0:
            Fl, _, nr = _rank_rows(rank - 1)
            assert torch.equal(r["local"][pos:pos + nr], Fl[-nr:])          # left halo = left neighbour's suffix
            pos += nr
        if rank < world - 1:
            Fr, nl, _ = _rank_rows(rank + 1)
            assert torch.equal(r["local"][pos:pos + nl], Fr[:nl])           # right halo = right neighbour's prefix
            pos += nl
        assert pos == r["local"].shape[0]
        want = torch.cat([_rank_rows(q)[0][:, 0].long() for q in range(world)])
        assert torch.equal(r["all"], want) and r["counts"] == [4 + q for q in range(world)]
        assert r["total"] == sum(4 + q for q in range(world))
        assert r["empty"].tolist() == [7] and r["ecounts"] == [1] + [0] * (world - 1)


def _worker(rank, world, port, out):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import pcc_b200
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r = _scenario(pcc_b200.slab.DistComm())
        torch.save(r, f"{out}.{rank}")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_dist_comm_gloo(tmp_path, world):
    out = str(tmp_path / "r")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    _check([torch.load(f"{out}.{r}") for r in range(world)], world)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_thread_comm(world):
    import pcc_b200
    comms = pcc_b200.slab.ThreadComm.make(world)
    results = [None] * world

    def run(r):
        results[r] = _scenario(comms[r])

    ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join(60) for t in ts]
    assert all(r is not None for r in results)
    _check(results, world)


def test_slab_bounds():
    import pcc_b200
    sb = pcc_b200.slab.slab_bounds
    assert sb(10, 1000, 4) == [0, 256, 512, 768, 1024]
    assert sb(0, 255, 8) == [0, 32, 64, 96, 128, 160, 192, 224, 256]
    b = sb(100, 900, 3, [0, 0, 0, 5, 5, 5, 100] + [1] * 19)
    assert b[0] == 96 and b[-1] == 928 and all(v % 32 == 0 for v in b) and b == sorted(set(b))
    with pytest.raises(pcc_b200.PccError):
        sb(0, 40, 4)
    # every slab is non-empty in cells even when all points sit in one cell
    b = sb(0, 255, 4, [1000, 0, 0, 0, 0, 0, 0, 0])
    assert len(set(b)) == 5
