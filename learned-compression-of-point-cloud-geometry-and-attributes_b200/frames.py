"""Frame-parallel sharding (BASELINE.json configs[3]).

Frames of a sequence are independent (evaluate.py:95-110 codes one frame at a time; weights are
read-only), so the path shards by frame with NO data-path collective: rank r codes frames
r, r + W, r + 2W, ...  The only communication is the reduction of the timing counters at the end
(max over ranks of the device time, sum of the points coded), over whatever backend the process
group was created with (NCCL on the GPU box, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def frames_of_rank(n_frames, rank, world):
    """Indices of the frames rank `rank` codes (round-robin, so every rank gets ceil or floor of n/W)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, n_frames, world))


def reduce_throughput(seconds, points, device=None):
    """-> (max over ranks of each entry of `seconds`, total points).  No-op without a process group."""
    seconds = [float(s) for s in (seconds if isinstance(seconds, (list, tuple)) else [seconds])]
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return seconds, int(points)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor(seconds, dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n = torch.tensor([int(points)], dtype=torch.int64, device=device)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return t.tolist(), int(n.item())
