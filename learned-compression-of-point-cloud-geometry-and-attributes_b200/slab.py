"""Spatial partition of ONE large scene across GPUs (BASELINE.json configs[4]).

The scene's bounding box is cut into x-slabs whose boundaries are multiples of 32 voxels, so a
voxel and all its stride-2/4/8/16/32 parents fall into the same slab.  Rank r owns the voxels with
x in [x0_r, x1_r) at every level.  A 3x3x3 convolution at tensor stride ts reads neighbours at
+-ts, so besides its owned rows a rank keeps, per level, a **one-voxel ring** of halo rows: the
left neighbour's voxels with x in [x0 - ts, x0) and the right neighbour's with x in [x1, x1 + ts).
Coordinates are exchanged once per map, features once per layer (`SlabMap.sync`), by point-to-point
send/recv with the two neighbours (NCCL over NVLink on the GPU box).

Row layout of a level on a rank: [owned rows, canonical order | left halo | right halo].  Owned
rows are in canonical (x-major) order, so the slab's boundary voxels are a prefix and a suffix of
the owned rows — what a rank sends is two contiguous slices, no gather — and the concatenation of
the ranks' owned rows in rank order IS the scene's canonical order.  That makes the two global
steps simple and exact:

* top-k pruning (model/blocks.py:130-150, k is scene-wide): the ranks all-gather channel 0 of the
  candidate logits (4 B per candidate), run the same radix select on the concatenation and keep the
  selected rows that fall into their own range — identical set and tie-breaking to the
  unpartitioned codec;
* entropy coding: the stride-8 latent (18 k voxels per 800 k points) is all-gathered and coded on
  every rank's copy identically (rank 0's strings are the result), so the bitstream is the single
  sequential rANS stream of the reference, byte-identical to the one-GPU result.

Every convolution output of an owned row sums the same neighbours in the same order as on one GPU
(halo rows are bit copies), hence the partitioned codec is bit-exact against the unpartitioned one
(tests/test_gpu_slab.py; tests/test_slab_gloo.py for the exchange logic on CPU).
"""
import threading

import torch

from . import ops
from ._lib import PccError

ALIGN = 32  # slab boundaries are multiples of this (deepest tensor stride on the path)


# ------------------------------------------------------------------------------- communicators
class DistComm:
    """torch.distributed process group (NCCL on GPUs, gloo on CPU tensors in the tests)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def exchange(self, left_send, right_send, out_left, out_right):
        """Send `left_send` to rank-1 and `right_send` to rank+1; receive into out_left (from rank-1)
        and out_right (from rank+1).  Tensors with zero rows are skipped on both sides."""
        d, ops_ = self.dist, []
        if self.rank > 0:
            if left_send.shape[0]:
                ops_.append(d.P2POp(d.isend, left_send, self.rank - 1, self.group))
            if out_left.shape[0]:
                ops_.append(d.P2POp(d.irecv, out_left, self.rank - 1, self.group))
        if self.rank < self.world - 1:
            if right_send.shape[0]:
                ops_.append(d.P2POp(d.isend, right_send, self.rank + 1, self.group))
            if out_right.shape[0]:
                ops_.append(d.P2POp(d.irecv, out_right, self.rank + 1, self.group))
        if ops_:
            for w in d.batch_isend_irecv(ops_):
                w.wait()

    def all_gather_rows(self, t):
        """-> (concatenation of every rank's rows in rank order, list of row counts)."""
        d = self.dist
        n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
        counts = [torch.empty_like(n) for _ in range(self.world)]
        d.all_gather(counts, n, group=self.group)
        counts = [int(c.item()) for c in counts]
        m = max(counts)
        pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[:t.shape[0]] = t
        parts = [torch.empty_like(pad) for _ in range(self.world)]
        d.all_gather(parts, pad, group=self.group)
        return torch.cat([p[:c] for p, c in zip(parts, counts)], dim=0), counts

    def all_reduce_sum(self, value, device):
        t = torch.tensor([int(value)], dtype=torch.int64, device=device)
        self.dist.all_reduce(t, group=self.group)
        return int(t.item())


class ThreadComm:
    """`world` virtual ranks as threads of one process sharing one GPU (tests, and a way to check a
    partition without a multi-GPU box).  Every collective is a rendezvous on a barrier; all threads
    enqueue on the same CUDA stream, so a tensor posted before the barrier is safe to read after it."""

    class _Shared:
        def __init__(self, world):
            self.world = world
            self.barrier = threading.Barrier(world)
            self.slots = [None] * world

    def __init__(self, shared, rank):
        self.s, self.rank, self.world = shared, rank, shared.world

    @staticmethod
    def make(world):
        shared = ThreadComm._Shared(world)
        return [ThreadComm(shared, r) for r in range(world)]

    @staticmethod
    def run(world, fn):
        """Run fn(comm) on `world` virtual ranks; returns the list of results.  An exception on one
        rank breaks the barrier (so the others do not wait for ever) and is re-raised."""
        comms = ThreadComm.make(world)
        results, errors = [None] * world, []

        def body(r):
            try:
                results[r] = fn(comms[r])
            except threading.BrokenBarrierError:
                pass
            except BaseException as e:  # noqa: BLE001 — re-raised below
                errors.append(e)
                comms[r].s.barrier.abort()

        threads = [threading.Thread(target=body, args=(r,)) for r in range(world)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return results

    def _rendezvous(self, item):
        self.s.slots[self.rank] = item
        self.s.barrier.wait()
        snapshot = list(self.s.slots)
        self.s.barrier.wait()
        return snapshot

    def exchange(self, left_send, right_send, out_left, out_right):
        posted = self._rendezvous((left_send, right_send))
        if self.rank > 0 and out_left.shape[0]:
            out_left.copy_(posted[self.rank - 1][1])
        if self.rank < self.world - 1 and out_right.shape[0]:
            out_right.copy_(posted[self.rank + 1][0])
        self.s.barrier.wait()  # the senders' buffers may be overwritten only after every copy is enqueued

    def all_gather_rows(self, t):
        posted = self._rendezvous(t)
        out = torch.cat(list(posted), dim=0)
        self.s.barrier.wait()
        return out, [int(p.shape[0]) for p in posted]

    def all_reduce_sum(self, value, device):
        return sum(int(v) for v in self._rendezvous(int(value)))


# ------------------------------------------------------------------------------- partition
def slab_bounds(x_min, x_max, world, counts_per_cell=None):
    """W + 1 ascending boundaries, multiples of ALIGN, covering [x_min, x_max].  With
    counts_per_cell (points per ALIGN-wide cell, starting at floor(x_min / ALIGN)) the cut
    balances the point counts; otherwise the cells are split evenly."""
    lo = (int(x_min) // ALIGN) * ALIGN
    hi = (int(x_max) // ALIGN + 1) * ALIGN
    cells = (hi - lo) // ALIGN
    if cells < world:
        raise PccError(f"scene spans {cells} cells of {ALIGN} voxels along x: too narrow for {world} slabs")
    if counts_per_cell is None:
        cuts = [round(cells * r / world) for r in range(world + 1)]
    else:
        c = [int(v) for v in counts_per_cell][:cells]
        c += [0] * (cells - len(c))
        total, cum, acc = sum(c), [], 0
        for v in c:
            acc += v
            cum.append(acc)
        cuts = [0]
        for r in range(1, world):
            i = next((j for j, v in enumerate(cum) if v >= total * r / world), cells - 1)
            cuts.append(min(max(i + 1, cuts[-1] + 1), cells - (world - r)))
        cuts.append(cells)
    for a, b in zip(cuts[:-1], cuts[1:]):
        if b <= a:
            raise PccError("empty slab in the partition")
    return [lo + ALIGN * c for c in cuts]


class Slab:
    """Rank `comm.rank`'s slab of a partition given by `bounds` (len world + 1)."""

    def __init__(self, comm, bounds):
        if len(bounds) != comm.world + 1 or any(b % ALIGN for b in bounds):
            raise PccError("bounds must hold world + 1 multiples of 32")
        self.comm, self.bounds = comm, [int(b) for b in bounds]
        # the outer slabs are open-ended: generative candidates may step outside the input's bounding box
        self.x0 = -32768 if comm.rank == 0 else self.bounds[comm.rank]
        self.x1 = 32768 if comm.rank == comm.world - 1 else self.bounds[comm.rank + 1]

    @staticmethod
    def _xkey(x):
        """Smallest key of batch 0 with the given x (keys order lexicographically by b, x, y, z)."""
        x = max(-32768, min(32768, int(x)))
        return (x + 32768) << 32

    def row_range(self, keys, xa, xb):
        """[first, last) rows of the sorted batch-0 `keys` with xa <= x < xb."""
        q = torch.tensor([self._xkey(xa), self._xkey(xb)], dtype=torch.int64, device=keys.device)
        a, b = torch.searchsorted(keys, q).tolist()
        return a, b


class SlabMap:
    """One level of one rank: owned coordinates + halo ring.  Quacks like ops.CoordMap for the
    codec (`keys`, `n`, `ts`, `coords`, `kernel_map`, `stride2`, `expand`, `lookup`); feature
    tensors on it have `n_local` rows ([owned | left halo | right halo])."""

    def __init__(self, own_keys, ts, slab):
        self.slab, self.ts = slab, int(ts)
        self.own = ops.CoordMap(own_keys, ts, True)
        n = self.own.n
        if n and (int((own_keys[-1] >> 48).item()) != 0):
            raise PccError("the slab partition handles one scene (batch index 0)")
        if n:
            a, b = slab.row_range(own_keys, slab.x0, slab.x1)
            if a != 0 or b != n:
                raise PccError("owned coordinates outside the rank's slab")
        # boundary voxels: prefix x < x0 + ts (left neighbour's right halo), suffix x >= x1 - ts
        self.n_send_left = slab.row_range(own_keys, slab.x0, slab.x0 + self.ts)[1] if n else 0
        self.n_send_right = n - slab.row_range(own_keys, slab.x1 - self.ts, slab.x1)[0] if n else 0
        comm, dev = slab.comm, own_keys.device
        cnt = torch.tensor([self.n_send_left, self.n_send_right], dtype=torch.int64, device=dev)
        got = torch.zeros(2, dtype=torch.int64, device=dev)
        comm.exchange(cnt[0:1], cnt[1:2], got[0:1], got[1:2])
        self.n_halo_left, self.n_halo_right = got.tolist()
        local = torch.empty(n + self.n_halo_left + self.n_halo_right, dtype=torch.int64, device=dev)
        local[:n] = own_keys
        self.sync(local)
        self.full = ops.CoordMap(local, ts, False)
        self._derived = {}
        self._coords = None

    # ---- CoordMap surface
    @property
    def keys(self):
        return self.own.keys

    @property
    def n(self):
        return self.own.n

    @property
    def n_local(self):
        return self.full.n

    @property
    def device(self):
        return self.own.device

    @property
    def coords(self):
        return self.own.coords

    def lookup(self, q_keys):
        return self.full.lookup(q_keys)

    def kernel_map(self, out_map, kernel_size, transpose=False, groups=None):
        """Inputs: every local row (owned + halo); outputs: out_map's owned rows."""
        out = out_map.own if isinstance(out_map, SlabMap) else out_map
        return self.full.kernel_map(out, kernel_size, transpose=transpose, groups=groups)

    def stride2(self):
        if "s2" not in self._derived:
            self._derived["s2"] = SlabMap(self.own.stride2().keys, self.ts * 2, self.slab)
        return self._derived["s2"]

    def expand(self, kernel_size):
        """Generative children of the local parents that fall into this slab."""
        key = ("ex", kernel_size)
        if key not in self._derived:
            cand = self.full.expand(kernel_size).keys
            a, b = self.slab.row_range(cand, self.slab.x0, self.slab.x1)
            self._derived[key] = SlabMap(cand[a:b].clone(), self.ts // 2, self.slab)
        return self._derived[key]

    # ---- slab-specific
    def sync(self, F):
        """Fill the halo rows of F ([n_local, ...]) with the neighbours' boundary rows."""
        n = self.own.n
        hl, hr = self.n_halo_left, self.n_halo_right
        self.slab.comm.exchange(F[:self.n_send_left], F[n - self.n_send_right:n],
                                F[n:n + hl], F[n + hl:n + hl + hr])
        return F

    def lift(self, F_own):
        """[n, C] features of the owned rows -> [n_local, C] with the halo rows filled."""
        F = torch.empty((self.n_local,) + tuple(F_own.shape[1:]), dtype=F_own.dtype, device=F_own.device)
        F[:self.own.n] = F_own
        return self.sync(F)

    def conv(self, feats, out_map, layer, transpose, epilogue):
        """One convolution into out_map (a SlabMap): owned rows computed, halo rows exchanged."""
        kmap = self.kernel_map(out_map, layer.kernel_size, transpose=transpose)
        F = torch.empty((out_map.n_local, layer.out_channels), dtype=torch.float32, device=feats.device)
        ops.conv(feats, kmap, layer.kernel, layer.bias, out=F[:out_map.n], **epilogue)
        return out_map.sync(F)

    def count_total(self):
        return self.slab.comm.all_reduce_sum(self.own.n, self.device)

    def topk(self, logits, k):
        """Scene-wide top-k over channel 0 of the owned candidates' logits -> selected owned rows
        (ascending int32), exactly the rows the unpartitioned select keeps."""
        mine = logits[:self.own.n, 0].contiguous()
        allv, counts = self.slab.comm.all_gather_rows(mine)
        sel = ops.topk_rows(allv.unsqueeze(1), int(k)).long()
        off = sum(counts[:self.slab.comm.rank])
        a, b = torch.searchsorted(sel, torch.tensor([off, off + self.own.n], device=sel.device)).tolist()
        return (sel[a:b] - off).to(torch.int32)

    def select(self, rows):
        return SlabMap(ops.take_keys(self.own.keys, rows), self.ts, self.slab)
