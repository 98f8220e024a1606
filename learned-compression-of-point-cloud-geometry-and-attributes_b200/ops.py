"""Functional layer over the C-ABI (include/pcc_b200.h): torch tensors in, raw pointers out.

torch is plumbing here (device memory, streams); every computation is a ``pcc_*`` call into
libpcc_b200.so.  There is no CPU fallback: tensors must live on a CUDA device.
"""
import ctypes
import os
import threading
import weakref

import torch

from . import _lib
from ._lib import ConvArgs, WgradArgs, check

TILE_M = 128
ACT_NONE, ACT_RELU, ACT_LEAKY = 0, 1, 2
MODE_EXACT, MODE_TF32 = 0, 1
LEAKY_SLOPE = 0.01  # ME.MinkowskiLeakyReLU default

_launches = 0  # kernels enqueued through this module (bench.py reports it as gpu_launches)
_count_lock = threading.Lock()
cache_lock = threading.RLock()  # guards the creation of device tensors cached across frames / threads


def launches():
    return _launches


def _count(n=1):
    global _launches
    with _count_lock:  # frames are launched from several threads (pipeline.FramePipeline)
        _launches += n


def publish(*tensors):
    """A device tensor about to be cached and read from OTHER streams (frame threads each have their own):
    wait until the current stream has produced it.  One-off per cache entry."""
    torch.cuda.current_stream().synchronize()
    return tensors[0] if len(tensors) == 1 else tensors


def _L():
    return _lib.load()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    if t is None:
        return None
    return ctypes.c_void_p(t.data_ptr())


def _need_cuda(t, name="tensor"):
    if not t.is_cuda:
        raise _lib.PccError(f"{name} must be a CUDA tensor: the product path has no CPU fallback")


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


# ---------------------------------------------------------------------------- coordinates
def pack_coords(coords):
    """[N,4] int32 (b,x,y,z) -> int64 tensor holding the uint64 keys.  Raises on out-of-range."""
    _need_cuda(coords, "coords")
    coords = coords.to(torch.int32).contiguous()
    n = coords.shape[0]
    keys = torch.empty(n, dtype=torch.int64, device=coords.device)
    status = torch.zeros(1, dtype=torch.int32, device=coords.device)
    check(_L().pcc_coords_pack(_ptr(coords), n, _ptr(keys), _ptr(status), _stream()), "pcc_coords_pack")
    _count()
    return keys, status


def unpack_keys(keys):
    n = keys.shape[0]
    coords = torch.empty((n, 4), dtype=torch.int32, device=keys.device)
    check(_L().pcc_coords_unpack(_ptr(keys), n, _ptr(coords), _stream()), "pcc_coords_unpack")
    _count()
    return coords


def sort_keys(keys, want_perm=True):
    """Canonical (lexicographic b,x,y,z) order.  Returns (sorted_keys, perm int32 or None)."""
    n = keys.shape[0]
    out = torch.empty_like(keys)
    perm = torch.empty(n, dtype=torch.int32, device=keys.device) if want_perm else None
    if n == 0:
        return out, perm
    wsb = _L().pcc_sort_ws_bytes(n)
    ws = _ws(wsb, keys.device)
    check(_L().pcc_sort_keys(_ptr(keys), n, _ptr(out), _ptr(perm), _ptr(ws), wsb, _stream()), "pcc_sort_keys")
    _count(4)
    return out, perm


class KernelMap:
    """nbr [K, stride] int32 (offset-major, indexed by TILE ROW of the offset's group), tapmask
    [n_groups * ceil(n_out/128)] uint32 (as int32), order [n_groups, n_out] int32: tile row -> output
    row (None = identity).  n_groups > 1: the K offsets are cut into consecutive ranges, each with its
    own tile-row permutation (include/pcc_b200.h, pcc_kernel_map)."""
    __slots__ = ("nbr", "tapmask", "K", "n_out", "stride", "order", "n_groups", "tile_order", "_pairs_cache")

    def __init__(self, nbr, tapmask, K, n_out, stride, order=None, n_groups=1, tile_order=None):
        self.nbr, self.tapmask, self.K, self.n_out, self.stride = nbr, tapmask, K, n_out, stride
        self.order, self.n_groups, self.tile_order = order, n_groups, tile_order

    def group_ranges(self):
        """[(k_lo, k_hi)] of the offset groups."""
        per = (self.K + self.n_groups - 1) // self.n_groups
        return [(g * per, min((g + 1) * per, self.K)) for g in range(self.n_groups)]

    def natural_nbr(self):
        """nbr re-indexed by output row ([K, n_out]); what the oracle's kernel map holds."""
        nbr = self.nbr[:, :self.n_out]
        if self.order is None:
            return nbr
        out = torch.empty_like(nbr)
        for g, (lo, hi) in enumerate(self.group_ranges()):
            out[lo:hi, self.order.view(self.n_groups, -1)[g].long()] = nbr[lo:hi]
        return out


class CoordMap:
    """One coordinate map: unique keys in row order + (lazy) hash table + cached derived maps.

    Plays the role of one entry of ME's coordinate manager; maps derived from it (stride-2
    children, generative children, kernel maps) are cached on the object, which is what lets
    layers that share coordinates share maps.
    """

    def __init__(self, keys, ts, is_sorted, n_batches=None):
        self.keys = keys
        self.ts = int(ts)
        self.is_sorted = bool(is_sorted)
        # upper bound of (largest batch id + 1) when known on the host: read once where the coordinates
        # enter (canonical_map) and inherited by every derived map, so that per-batch logic (point
        # counts, per-batch top-k) never has to synchronise the stream to ask the device
        self.n_batches = None if n_batches is None else int(n_batches)
        self._hash = None
        self._coords = None
        self._derived = {}
        self._kmaps = {}

    @property
    def n(self):
        return self.keys.shape[0]

    @property
    def device(self):
        return self.keys.device

    @property
    def coords(self):
        if self._coords is None:
            self._coords = unpack_keys(self.keys)
        return self._coords

    def hash(self):
        if self._hash is None:
            cap = _L().pcc_hash_capacity(self.n)
            tkeys = torch.empty(cap, dtype=torch.int64, device=self.device)
            tvals = torch.empty(cap, dtype=torch.int32, device=self.device)
            check(_L().pcc_hash_build(_ptr(self.keys), self.n, _ptr(tkeys), _ptr(tvals), cap, _stream()),
                  "pcc_hash_build")
            _count(2)
            self._hash = (tkeys, tvals, cap)
        return self._hash

    def lookup(self, q_keys):
        """rows of q_keys in this map (-1 when absent)."""
        nq = q_keys.shape[0]
        rows = torch.empty(nq, dtype=torch.int32, device=self.device)
        if nq:
            tk, tv, cap = self.hash()
            check(_L().pcc_hash_probe(_ptr(tk), _ptr(tv), cap, _ptr(q_keys), nq, _ptr(rows), _stream()),
                  "pcc_hash_probe")
            _count()
        return rows

    def stride2(self):
        """Output map of a stride-2 convolution (sorted)."""
        if "s2" not in self._derived:
            n = self.n
            out = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
            n_out = ctypes.c_int64(0)
            wsb = _L().pcc_stride_ws_bytes(n)
            ws = _ws(wsb, self.device)
            check(_L().pcc_coords_stride(_ptr(self.keys), n, self.ts * 2, _ptr(out), ctypes.byref(n_out),
                                         _ptr(ws), wsb, _stream()), "pcc_coords_stride")
            _count(8)
            self._derived["s2"] = CoordMap(out[:n_out.value].clone() if n_out.value < n else out[:n_out.value],
                                           self.ts * 2, True, self.n_batches)
        return self._derived["s2"]

    def expand(self, kernel_size):
        """Generative output map of a stride-2 transposed convolution (sorted)."""
        key = ("ex", kernel_size)
        if key not in self._derived:
            n = self.n
            K = kernel_size ** 3
            out = torch.empty(max(n * K, 1), dtype=torch.int64, device=self.device)
            n_out = ctypes.c_int64(0)
            wsb = _L().pcc_expand_ws_bytes(n, kernel_size)
            ws = _ws(wsb, self.device)
            tk, tv, cap = self.hash()
            check(_L().pcc_coords_expand(_ptr(tk), _ptr(tv), cap, _ptr(self.keys), n, kernel_size, self.ts,
                                         _ptr(out), ctypes.byref(n_out), _ptr(ws), wsb, _stream()),
                  "pcc_coords_expand")
            _count(8)
            self._derived[key] = CoordMap(out[:n_out.value].clone(), self.ts // 2, True, self.n_batches)
        return self._derived[key]

    def kernel_map(self, out_map, kernel_size, transpose=False, groups=None):
        """Kernel map from this (input) map to out_map.  conv: targets c_o + d*ts_in;
        transpose: targets q - d*ts_out, probed only when aligned to ts_in.  groups: number of
        offset groups (None = tap_groups(): by mode and size)."""
        K = kernel_size ** 3
        n_out = out_map.n
        seg = tile_segment_rows()
        ng = tap_groups(K, n_out) if groups is None else max(0, min(int(groups), K))  # 0 = measured choice
        key = (id(out_map), kernel_size, transpose, ng, seg)
        km = self._kmaps.get(key)
        if km is not None and km[0]() is out_map:
            return km[1]
        stride = max(n_out, 1)
        nbr = torch.empty((K, stride), dtype=torch.int32, device=self.device)
        cap_g = ng if ng else AUTO_GROUPS
        tiles = max((n_out + TILE_M - 1) // TILE_M, 1)
        tapmask = torch.empty(cap_g * tiles, dtype=torch.int32, device=self.device)
        order = tile_order = None
        if n_out:
            tk, tv, cap = self.hash()
            step = -out_map.ts if transpose else self.ts
            align = self.ts if transpose else 1
            ws, wsb = None, 0
            if not (K > 1 and n_out > TILE_M and seg):
                ng = 1
            else:
                order = torch.empty(cap_g * n_out, dtype=torch.int32, device=self.device)
                tile_order = torch.empty(cap_g * tiles, dtype=torch.int32, device=self.device)
                wsb = _L().pcc_kernel_map_ws_bytes(n_out)
                ws = _ws(wsb, self.device)
            used = ctypes.c_int32(1)
            check(_L().pcc_kernel_map(_ptr(tk), _ptr(tv), cap, _ptr(out_map.keys), n_out, kernel_size, step,
                                      align, _ptr(nbr), stride, _ptr(tapmask), _ptr(order), _ptr(tile_order), seg, ng,
                                      ctypes.byref(used), _ptr(ws), wsb, _stream()), "pcc_kernel_map")
            if ng == 0:
                _count(6)
            ng = used.value
            if order is not None:
                order, tapmask, tile_order = order[:ng * n_out], tapmask[:ng * tiles], tile_order[:ng * tiles]
            _count(1 if order is None else 1 + 7 * ng)
        else:
            ng = 1
        kmap = KernelMap(nbr, tapmask, K, n_out, stride, order, ng, tile_order)
        # weak reference: out_map is often this map itself (stride-1 layers) or a map that caches the
        # transposed direction back to this one; a strong reference would make the pair collectable
        # only by the cycle collector and keep its kernel maps (K x n_out x 4 bytes) alive until then
        self._kmaps[key] = (weakref.ref(out_map), kmap)
        return kmap


# Rows per segment of the tile-row sort (0 = keep lexicographic tiles).  None = by mode: the exact
# (FP32-pipe-bound) path sorts whole maps; the tensor-core path keeps a tile's gathers inside an
# L2-sized segment of the map.
_tile_seg_rows = int(os.environ["PCC_TILE_SEG_ROWS"]) if "PCC_TILE_SEG_ROWS" in os.environ else None  # A/B override


def set_tile_segment_rows(n):
    """Segment length of the kernel map's tile-row sort (power of two >= 128; 0 disables it).
    Affects only maps built afterwards; results are independent of it."""
    global _tile_seg_rows
    _tile_seg_rows = None if n is None else int(n)


def tile_segment_rows():
    """The segment length in effect for maps built now."""
    if _tile_seg_rows is not None:
        return _tile_seg_rows
    return (1 << 30) if _default_mode == MODE_EXACT else (1 << 17)


# Offset groups of the kernel maps (pcc_kernel_map n_groups).  None = by mode: the exact path lets the
# library measure each 27-offset map (n_groups = 0): maps whose mask-sorted tiles use >= 90 % of their MACs, and
# maps too small to fill the GPU twice, keep one group; down to 70 % the 27 offsets are cut into 2 ranges, below
# that (e.g. the decoder's pruned sets under random-init weights, 40-60 %) into 3, each range with its own
# mask-sorted tile rows (89-99 %), for one or two extra chained launches per convolution (coords.cu has the
# measurements behind the rule).  The tensor-core kernel takes single-group maps only.
_tap_groups = int(os.environ["PCC_TAP_GROUPS"]) if "PCC_TAP_GROUPS" in os.environ else None
AUTO_GROUPS = 3


def set_tap_groups(n):
    """Force the number of offset groups of 27-offset kernel maps built afterwards (None = automatic).
    Results are independent of it."""
    global _tap_groups
    _tap_groups = None if n is None else int(n)


def tap_groups(K, n_out):
    if K != 27:
        return 1
    if _tap_groups is not None:
        return max(0, min(_tap_groups, K))
    return 0 if (_default_mode == MODE_EXACT and n_out > TILE_M) else 1


def make_map(coords_or_keys, ts=1, assume_sorted=False):
    """CoordMap from [N,4] int coordinates (or packed keys), rows kept in the given order."""
    if coords_or_keys.dim() == 2:
        keys, status = pack_coords(coords_or_keys)
        if int(status.item()) != 0:
            raise _lib.PccError("coordinate out of the packable range (|x|,|y|,|z| < 32768, 0 <= b < 65536)")
    else:
        keys = coords_or_keys
    return CoordMap(keys, ts, assume_sorted)


def canonical_map(coords, ts=1):
    """Sort coordinates into canonical order.  Returns (CoordMap sorted, perm int64) and checks
    uniqueness (one host sync)."""
    keys, status = pack_coords(coords)
    skeys, perm = sort_keys(keys)
    dup = (skeys[1:] == skeys[:-1]).any() if skeys.shape[0] > 1 else torch.zeros((), dtype=torch.bool, device=keys.device)
    last_batch = (skeys[-1] >> 48) if skeys.shape[0] else torch.zeros((), dtype=torch.int64, device=keys.device)
    flags = torch.stack([(status[0] != 0).long(), dup.reshape(()).long(), last_batch.reshape(()).long()]).tolist()
    if flags[0]:
        raise _lib.PccError("coordinate out of the packable range (|x|,|y|,|z| < 32768, 0 <= b < 65536)")
    if flags[1]:
        raise _lib.PccError("duplicate coordinates")
    return CoordMap(skeys, ts, True, flags[2] + 1 if skeys.shape[0] else 0), perm.long()


# ---------------------------------------------------------------------------- convolution
_default_mode = MODE_EXACT
_profile = None  # list of per-launch records while conv profiling is on (bench.py roofline pass)


def start_conv_profile():
    """Record CUDA events (on the launching stream) around every pcc_conv_fwd launch."""
    global _profile
    _profile = []


def stop_conv_profile():
    """-> list of dicts: cin, cout, K, n_in, n_out, pairs, mode, ms (after a device sync)."""
    global _profile
    recs, _profile = _profile, None
    torch.cuda.synchronize()
    out = []
    for r in recs or []:
        r = dict(r)
        r["ms"] = r.pop("ev0").elapsed_time(r.pop("ev1"))
        out.append(r)
    return out


def _pairs(kmap):
    if not hasattr(kmap, "_pairs_cache"):
        kmap._pairs_cache = int((kmap.nbr[:, :kmap.n_out] >= 0).sum().item()) if kmap.n_out else 0
    return kmap._pairs_cache


def set_default_mode(mode):
    """PCC_MODE_EXACT (bit-exact vs oracle) or PCC_MODE_TF32 (tcgen05 where the layer allows)."""
    global _default_mode
    _default_mode = int(mode)


def default_mode():
    return _default_mode


def tf32_eligible(cin, cout):
    return cin % 32 == 0 and cout % 32 == 0 and 32 <= cout <= 256


_wt_cache = {}


def transposed_weight(weight):
    """[K, cin, cout] -> cached [K, cout, cin] copy for the tensor-core path (pcc_weight_transpose).
    The cache is keyed on the tensor OBJECT (weak reference) and its in-place version counter, so a
    modified or re-created parameter is re-laid-out and a recycled address can never alias."""
    import weakref
    key = id(weight)
    ver = (weight._version, tuple(weight.shape), str(weight.device), weight.data_ptr())
    hit = _wt_cache.get(key)
    if hit is not None and hit[0]() is weight and hit[1] == ver:
        return _wt_acquire(hit)
    with cache_lock:
        return _wt_acquire(_transposed_weight_build(weight, key, ver))


def _wt_acquire(hit):
    """The copy was produced on hit[4]; any other stream orders itself behind the producing kernel with an
    event wait (no host synchronisation: in training every weight is re-laid-out every step)."""
    cur = torch.cuda.current_stream()
    if cur.cuda_stream != hit[4]:
        cur.wait_event(hit[3])
        hit[2].record_stream(cur)
    return hit[2]


def _transposed_weight_build(weight, key, ver):
    import weakref
    hit = _wt_cache.get(key)
    if hit is not None and hit[0]() is weight and hit[1] == ver:
        return hit
    if len(_wt_cache) > 256:
        for k in [k for k, v in _wt_cache.items() if v[0]() is None]:
            del _wt_cache[k]
    w = weight.detach()
    w = (w if w.dim() == 3 else w.unsqueeze(0)).contiguous()
    K, cin, cout = w.shape
    wt = torch.empty((K, cout, cin), dtype=torch.float32, device=w.device)
    check(_L().pcc_weight_transpose(_ptr(w), K, cin, cout, _ptr(wt), _stream()), "pcc_weight_transpose")
    _count()
    ev = torch.cuda.Event()
    ev.record()
    hit = (weakref.ref(weight), ver, wt, ev, torch.cuda.current_stream().cuda_stream)
    _wt_cache[key] = hit
    return hit


CONV_ROUND_OUT_TF32 = 1


def conv(feats, kmap, weight, bias=None, act=ACT_NONE, film=None, film_rows=None, residual=None,
         res_rows=None, out=None, mode=None, round_out=None, weight_t=None):
    """out[o] = epilogue(sum_k feats[nbr[k,o]] @ weight[k]).

    weight: [K, cin, cout] (or [cin, cout] for K == 1), bias: [1, cout] or [cout].
    film: [n_film, 2*cf] (beta | gamma), cf in {cout, 1}; residual: [n_res, cout].
    weight_t: the [K, cout, cin] layout of `weight` when the caller already holds it (the data gradient does:
    it is the forward weight); otherwise the tensor-core path takes it from the cache of transposed_weight.
    """
    _need_cuda(feats, "feats")
    feats = feats.contiguous()
    w = weight if weight.dim() == 3 else weight.unsqueeze(0)
    w = w.contiguous()
    K, cin, cout = w.shape
    if K != kmap.K or feats.shape[1] != cin:
        raise _lib.PccError(f"conv shape mismatch: feats {tuple(feats.shape)}, weight {tuple(w.shape)}, K {kmap.K}")
    n_out = kmap.n_out
    if out is None:
        out = torch.empty((n_out, cout), dtype=torch.float32, device=feats.device)
    if mode is None:
        mode = _default_mode
    if round_out is None:  # in tf32 mode every activation that can feed a tensor-core layer is stored tf32-rounded
        round_out = (mode == MODE_TF32 and cout % 32 == 0)
    if mode == MODE_TF32 and not tf32_eligible(cin, cout):
        mode = MODE_EXACT  # stated policy: small-channel layers always run the exact CUDA-core path
    a = ConvArgs()
    a.in_ = feats.data_ptr()
    a.n_in = feats.shape[0]
    a.cin, a.cout = cin, cout
    a.weight = w.data_ptr()
    if mode != MODE_TF32:
        wt = None
    elif weight_t is not None:
        wt = weight_t.contiguous()
        if tuple(wt.shape[-3:] if wt.dim() == 3 else (1,) + tuple(wt.shape)) != (K, cout, cin):
            raise _lib.PccError(f"weight_t {tuple(wt.shape)} is not the [K, cout, cin] layout of {tuple(w.shape)}")
    else:
        wt = transposed_weight(weight)
    a.weight_t = None if wt is None else wt.data_ptr()
    b = None if bias is None else bias.contiguous()
    a.bias = None if b is None else b.data_ptr()
    a.nbr = kmap.nbr.data_ptr()
    a.nbr_stride = kmap.stride
    a.tapmask = kmap.tapmask.data_ptr()
    a.K = K
    a.act = act
    a.slope = LEAKY_SLOPE
    keep = [feats, w, b]
    if film is not None:
        film = film.contiguous()
        cf = film.shape[1] // 2
        a.film_c = cf
        a.film = film.data_ptr()
        a.film_rows = None if film_rows is None else film_rows.data_ptr()
        keep += [film, film_rows]
    else:
        a.film_c = 0
    if residual is not None:
        residual = residual.contiguous()
        a.residual = residual.data_ptr()
        a.res_rows = None if res_rows is None else res_rows.data_ptr()
        keep += [residual, res_rows]
    a.out = out.data_ptr()
    a.n_out = n_out
    a.out_stride = out.stride(0) if n_out else cout
    a.mode = mode
    a.flags = CONV_ROUND_OUT_TF32 if round_out else 0
    a.order = None if kmap.order is None else kmap.order.data_ptr()
    a.n_groups = kmap.n_groups
    # longest-tiles-first launch order: only for the FP32-pipe-bound exact kernels; the tensor-core
    # kernel is gather-bound and wants neighbouring tiles (an L2-sized window of the map) in flight together
    a.tile_order = kmap.tile_order.data_ptr() if (kmap.tile_order is not None and mode == MODE_EXACT) else None
    if mode == MODE_TF32 and kmap.n_groups > 1:
        raise _lib.PccError("the tensor-core mode takes single-group kernel maps: build the map in tf32 mode "
                            "or with set_tap_groups(1)")
    if _profile is not None and n_out:
        rec = {"cin": cin, "cout": cout, "K": K, "n_in": int(feats.shape[0]), "n_out": int(n_out),
               "pairs": _pairs(kmap), "mode": int(mode), "ev0": torch.cuda.Event(enable_timing=True),
               "ev1": torch.cuda.Event(enable_timing=True)}
        rec["ev0"].record()
        check(_L().pcc_conv_fwd(ctypes.byref(a), _stream()), "pcc_conv_fwd")
        rec["ev1"].record()
        _profile.append(rec)
        _count(kmap.n_groups)
        return out
    check(_L().pcc_conv_fwd(ctypes.byref(a), _stream()), "pcc_conv_fwd")
    if n_out:
        _count(kmap.n_groups)
    return out


def conv_head0(feats, kmap, weight, bias=None):
    """Channel 0 of conv(feats, kmap, weight, bias) in factorised form -> [n_out, 1]
    (pcc_conv_head0: tolerance mode only; single-group kernel maps)."""
    _need_cuda(feats, "feats")
    feats = feats.contiguous()
    w = (weight if weight.dim() == 3 else weight.unsqueeze(0)).contiguous()
    K, cin, cout = w.shape
    if K != kmap.K or feats.shape[1] != cin or kmap.n_groups > 1:
        raise _lib.PccError("conv_head0: shape mismatch or multi-group kernel map")
    out = torch.empty((kmap.n_out, 1), dtype=torch.float32, device=feats.device)
    wsb = _L().pcc_conv_head0_ws_bytes(feats.shape[0])
    ws = _ws(wsb, feats.device)
    b = None if bias is None else bias.contiguous()
    check(_L().pcc_conv_head0(_ptr(feats), feats.shape[0], cin, _ptr(w), cout, _ptr(b), _ptr(kmap.nbr), kmap.stride,
                              _ptr(kmap.order), kmap.n_out, K, _ptr(out), _ptr(ws), wsb, _stream()), "pcc_conv_head0")
    if kmap.n_out:
        _count(2)
    return out


def gather_rows(feats, rows, add=None):
    """features_at_coordinates on-grid: out[i] = feats[rows[i]] (0 when rows[i] < 0) (+ add[i])."""
    feats = feats.contiguous()
    n = rows.shape[0] if rows is not None else feats.shape[0]
    c = feats.shape[1]
    out = torch.empty((n, c), dtype=torch.float32, device=feats.device)
    if add is not None:
        add = add.contiguous()
    check(_L().pcc_gather_rows(_ptr(feats), c, _ptr(rows), n, _ptr(add), _ptr(out), _stream()), "pcc_gather_rows")
    if n:
        _count()
    return out


def conv_wgrad(feats, gout, kmap, cin, cout, mode=MODE_EXACT):
    """Weight gradient of conv(): dW[k] = sum_j feats[nbr[k][j]]^T (x) gout[order[j]]  -> [K, cin, cout].
    Deterministic (fixed summation order).  SURVEY.md §8f row 4; ME ConvolutionBackward's grad_kernel."""
    _need_cuda(feats, "feats")
    feats, gout = feats.contiguous(), gout.contiguous()
    K = kmap.K
    if feats.shape[1] != cin or gout.shape[1] != cout or gout.shape[0] != kmap.n_out:
        raise _lib.PccError(f"conv_wgrad shape mismatch: feats {tuple(feats.shape)}, gout {tuple(gout.shape)}, "
                            f"n_out {kmap.n_out}")
    dw = torch.empty((K, cin, cout), dtype=torch.float32, device=feats.device)
    wsb = _L().pcc_conv_wgrad_ws_bytes(K, cin, cout)
    ws = _ws(wsb, feats.device)
    a = WgradArgs()
    a.in_, a.gout, a.cin, a.cout = feats.data_ptr(), gout.data_ptr(), cin, cout
    a.nbr, a.nbr_stride = kmap.nbr.data_ptr(), kmap.stride
    a.tapmask = kmap.tapmask.data_ptr()
    a.order = None if kmap.order is None else kmap.order.data_ptr()
    a.n_out, a.K, a.n_groups = kmap.n_out, K, kmap.n_groups
    a.dweight, a.ws, a.ws_bytes = dw.data_ptr(), ws.data_ptr(), wsb
    a.mode = int(mode)
    if _profile is not None and kmap.n_out:
        rec = {"cin": cin, "cout": cout, "K": K, "n_in": int(feats.shape[0]), "n_out": int(kmap.n_out),
               "pairs": _pairs(kmap), "mode": -1 if not (mode == MODE_TF32 and tf32_eligible(cin, cout)) else -2,
               "ev0": torch.cuda.Event(enable_timing=True),
               "ev1": torch.cuda.Event(enable_timing=True)}
        rec["ev0"].record()
        check(_L().pcc_conv_wgrad(ctypes.byref(a), _stream()), "pcc_conv_wgrad")
        rec["ev1"].record()
        _profile.append(rec)
    else:
        check(_L().pcc_conv_wgrad(ctypes.byref(a), _stream()), "pcc_conv_wgrad")
    _count(kmap.n_groups + 1)
    return dw


def col_sum(x):
    """[n, c] -> [c] column sums in a fixed order (the bias gradient; pcc_col_sum)."""
    _need_cuda(x, "x")
    x = x.contiguous()
    n, c = x.shape
    out = torch.empty((c,), dtype=torch.float32, device=x.device)
    wsb = _L().pcc_col_sum_ws_bytes(c)
    ws = _ws(wsb, x.device)
    check(_L().pcc_col_sum(_ptr(x), n, c, _ptr(out), _ptr(ws), wsb, _stream()), "pcc_col_sum")
    _count(2 if n else 0)
    return out


def scatter_rows_add(g, rows, n_dst):
    """Backward of gather_rows: out [n_dst, c] with out[rows[i]] += g[i] (rows < 0 skipped)."""
    g = g.contiguous()
    out = torch.zeros((n_dst, g.shape[1]), dtype=torch.float32, device=g.device)
    check(_L().pcc_scatter_rows_add(_ptr(g), g.shape[1], _ptr(rows), g.shape[0], _ptr(out), _stream()),
          "pcc_scatter_rows_add")
    if g.shape[0]:
        _count()
    return out


def gc_likelihood_train_fwd(y_noisy, params, rows, scale_bound, lik_bound):
    """-> likelihood [N, C] (row-major) of noise-quantised latents (GaussianConditional, training)."""
    n, c = y_noisy.shape
    lik = torch.empty((n, c), dtype=torch.float32, device=y_noisy.device)
    check(_L().pcc_gc_likelihood_train_fwd(_ptr(y_noisy.contiguous()), _ptr(params.contiguous()), _ptr(rows), n, c,
                                           scale_bound, lik_bound, _ptr(lik), _stream()),
          "pcc_gc_likelihood_train_fwd")
    _count()
    return lik


def gc_likelihood_train_bwd(y_noisy, params, rows, scale_bound, lik_bound, g_lik):
    """-> (g_y [N, C], g_param_rows [N, 2C])."""
    n, c = y_noisy.shape
    g_y = torch.empty((n, c), dtype=torch.float32, device=y_noisy.device)
    g_p = torch.empty((n, 2 * c), dtype=torch.float32, device=y_noisy.device)
    check(_L().pcc_gc_likelihood_train_bwd(_ptr(y_noisy.contiguous()), _ptr(params.contiguous()), _ptr(rows), n, c,
                                           scale_bound, lik_bound, _ptr(g_lik.contiguous()), _ptr(g_y), _ptr(g_p),
                                           _stream()), "pcc_gc_likelihood_train_bwd")
    _count()
    return g_y, g_p


def take_keys(keys, rows):
    out = torch.empty(rows.shape[0], dtype=torch.int64, device=keys.device)
    check(_L().pcc_take_keys(_ptr(keys), _ptr(rows), rows.shape[0], _ptr(out), _stream()), "pcc_take_keys")
    if rows.shape[0]:
        _count()
    return out


def topk_rows(logits, k, column=0):
    """Rows (ascending) of the k largest logits[:, column]; ties -> smaller row index."""
    assert logits.dim() == 2
    m = logits.shape[0]
    k = int(k)
    sel = torch.empty(k, dtype=torch.int32, device=logits.device)
    wsb = _L().pcc_topk_ws_bytes(m)
    ws = _ws(wsb, logits.device)
    base = logits.data_ptr() + 4 * column * logits.stride(1)
    check(_L().pcc_topk_select(ctypes.c_void_p(base), logits.stride(0), m, k, _ptr(sel), _ptr(ws), wsb, _stream()),
          "pcc_topk_select")
    if k:
        _count(17)
    return sel


# ---------------------------------------------------------------------------- entropy stage
def gc_quantize_index(y, params, rows, scale_table, scale_bound):
    """-> symbols [C,N] int32, indexes [C,N] int32 (channel-major, what the coder consumes)."""
    n, c = y.shape
    sym = torch.empty((c, n), dtype=torch.int32, device=y.device)
    idx = torch.empty((c, n), dtype=torch.int32, device=y.device)
    check(_L().pcc_gc_quantize_index(_ptr(y.contiguous()), _ptr(params.contiguous()), _ptr(rows), n, c,
                                     _ptr(scale_table), scale_table.shape[0], scale_bound, _ptr(sym), _ptr(idx),
                                     _stream()), "pcc_gc_quantize_index")
    _count()
    return sym, idx


def gc_build_indexes(params, rows, n, c, scale_table, scale_bound):
    idx = torch.empty((c, n), dtype=torch.int32, device=params.device)
    check(_L().pcc_gc_build_indexes(_ptr(params.contiguous()), _ptr(rows), n, c, _ptr(scale_table),
                                    scale_table.shape[0], scale_bound, _ptr(idx), _stream()), "pcc_gc_build_indexes")
    _count()
    return idx


def gc_dequantize(symbols, params, rows):
    c, n = symbols.shape
    y_hat = torch.empty((n, c), dtype=torch.float32, device=symbols.device)
    check(_L().pcc_gc_dequantize(_ptr(symbols.contiguous()), _ptr(params.contiguous()), _ptr(rows), n, c,
                                 _ptr(y_hat), _stream()), "pcc_gc_dequantize")
    _count()
    return y_hat


def gc_likelihood(y, params, rows, scale_bound, lik_bound):
    n, c = y.shape
    y_hat = torch.empty((n, c), dtype=torch.float32, device=y.device)
    lik = torch.empty((c, n), dtype=torch.float32, device=y.device)
    check(_L().pcc_gc_likelihood(_ptr(y.contiguous()), _ptr(params.contiguous()), _ptr(rows), n, c, scale_bound,
                                 lik_bound, _ptr(y_hat), _ptr(lik), _stream()), "pcc_gc_likelihood")
    _count()
    return y_hat, lik


def eb_quantize(z, medians):
    n, c = z.shape
    sym = torch.empty((c, n), dtype=torch.int32, device=z.device)
    z_hat = torch.empty((n, c), dtype=torch.float32, device=z.device)
    check(_L().pcc_eb_quantize(_ptr(z.contiguous()), _ptr(medians), n, c, _ptr(sym), _ptr(z_hat), _stream()),
          "pcc_eb_quantize")
    _count()
    return sym, z_hat


def eb_dequantize(symbols, medians):
    c, n = symbols.shape
    z_hat = torch.empty((n, c), dtype=torch.float32, device=symbols.device)
    check(_L().pcc_eb_dequantize(_ptr(symbols.contiguous()), _ptr(medians), n, c, _ptr(z_hat), _stream()),
          "pcc_eb_dequantize")
    _count()
    return z_hat


def eb_likelihood(z, medians, eb_params, lik_bound):
    n, c = z.shape
    z_hat = torch.empty((n, c), dtype=torch.float32, device=z.device)
    lik = torch.empty((c, n), dtype=torch.float32, device=z.device)
    check(_L().pcc_eb_likelihood(_ptr(z.contiguous()), _ptr(medians), _ptr(eb_params), n, c, lik_bound,
                                 _ptr(z_hat), _ptr(lik), _stream()), "pcc_eb_likelihood")
    _count()
    return z_hat, lik


def color_finalize(keys, feats):
    n, c = feats.shape
    out = torch.empty((n, 3 + c), dtype=torch.float32, device=feats.device)
    check(_L().pcc_color_finalize(_ptr(keys), _ptr(feats.contiguous()), n, c, _ptr(out), _stream()),
          "pcc_color_finalize")
    _count()
    return out


# ---------------------------------------------------------------------------- side channel
def octree_encode(keys, stride, depth):
    """Unique packed keys (batch 0, coordinates >= 0 on the `stride` grid) -> occupancy bytes."""
    _need_cuda(keys, "keys")
    n = keys.shape[0]
    if n == 0:
        return b""
    cap = ((n * depth + 3) // 4) * 4
    out = torch.empty(cap, dtype=torch.uint8, device=keys.device)
    wsb = _L().pcc_octree_ws_bytes(n)
    ws = _ws(wsb, keys.device)
    nbytes = ctypes.c_int64(0)
    check(_L().pcc_octree_encode(_ptr(keys), n, stride, depth, _ptr(out), cap, ctypes.byref(nbytes), _ptr(ws), wsb,
                                 _stream()), "pcc_octree_encode")
    _count(5)
    return out[:nbytes.value].cpu().numpy().tobytes()


def octree_decode(stream, stride, depth, device, cap=None):
    """Occupancy bytes -> packed keys in Morton order (int64 tensor on `device`)."""
    import numpy as np
    nbytes = len(stream)
    if nbytes == 0:
        return torch.empty(0, dtype=torch.int64, device=device)
    dev_bytes = torch.from_numpy(np.frombuffer(stream, dtype=np.uint8).copy()).to(device)
    cap = int(cap) if cap else 8 * nbytes
    keys = torch.empty(cap, dtype=torch.int64, device=device)
    wsb = _L().pcc_octree_ws_bytes(cap)
    ws = _ws(wsb, device)
    n_out = ctypes.c_int64(0)
    with torch.cuda.device(device):
        check(_L().pcc_octree_decode(_ptr(dev_bytes), nbytes, stride, depth, _ptr(keys), cap, ctypes.byref(n_out),
                                     _ptr(ws), wsb, _stream()), "pcc_octree_decode")
    _count(2)
    return keys[:n_out.value]


# ---------------------------------------------------------------------------- host coder
def _np_ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def rans_encode(symbols, indexes, cdf, cdf_length, offset):
    """numpy int32 arrays (host) -> bytes.  The reference's sequential CPU coder."""
    import numpy as np
    symbols = np.ascontiguousarray(symbols, dtype=np.int32).reshape(-1)
    indexes = np.ascontiguousarray(indexes, dtype=np.int32).reshape(-1)
    n = symbols.shape[0]
    cap = 4 * n + 64
    while True:
        out = np.empty(cap, dtype=np.uint8)
        nb = _L().pcc_rans_encode_with_indexes(_np_ptr(symbols), _np_ptr(indexes), n, _np_ptr(cdf), cdf.shape[1],
                                               _np_ptr(cdf_length), _np_ptr(offset), cdf.shape[0], _np_ptr(out), cap)
        if nb == -4:  # PCC_ERR_CAPACITY
            cap *= 4
            continue
        check(nb, "pcc_rans_encode_with_indexes")
        return out[:nb].tobytes()


def rans_decode(string, indexes, cdf, cdf_length, offset):
    import numpy as np
    indexes = np.ascontiguousarray(indexes, dtype=np.int32).reshape(-1)
    enc = np.frombuffer(string, dtype=np.uint8)
    out = np.empty(indexes.shape[0], dtype=np.int32)
    check(_L().pcc_rans_decode_with_indexes(_np_ptr(enc), enc.shape[0], _np_ptr(indexes), indexes.shape[0],
                                            _np_ptr(cdf), cdf.shape[1], _np_ptr(cdf_length), _np_ptr(offset),
                                            cdf.shape[0], _np_ptr(out)), "pcc_rans_decode_with_indexes")
    return out


def pmf_to_quantized_cdf(pmf, precision=16):
    import numpy as np
    pmf = np.ascontiguousarray(pmf, dtype=np.float32)
    cdf = np.empty(pmf.shape[0] + 1, dtype=np.uint32)
    check(_L().pcc_pmf_to_quantized_cdf(_np_ptr(pmf), pmf.shape[0], precision, _np_ptr(cdf)),
          "pcc_pmf_to_quantized_cdf")
    return cdf.astype(np.int32)


def gaussian_cdf_tables(scale_table, tail_mass=1e-9):
    """-> (cdf [n, stride] int32, cdf_length [n], offset [n]) numpy (host)."""
    import numpy as np
    st = np.ascontiguousarray(scale_table, dtype=np.float32)
    n = st.shape[0]
    stride = ctypes.c_int32(0)
    check(_L().pcc_gaussian_cdf_tables(_np_ptr(st), n, tail_mass, None, ctypes.byref(stride), None, None),
          "pcc_gaussian_cdf_tables")
    cdf = np.zeros((n, stride.value), dtype=np.int32)
    length = np.zeros(n, dtype=np.int32)
    offset = np.zeros(n, dtype=np.int32)
    check(_L().pcc_gaussian_cdf_tables(_np_ptr(st), n, tail_mass, _np_ptr(cdf), ctypes.byref(stride),
                                       _np_ptr(length), _np_ptr(offset)), "pcc_gaussian_cdf_tables")
    return cdf, length, offset


def factorized_cdf_tables(matrices, biases, factors, quantiles):
    """matrices/biases: 5 numpy float32 arrays, factors: 4, quantiles [C,1,3] -> tables (host)."""
    import numpy as np
    ms = [np.ascontiguousarray(m, dtype=np.float32) for m in matrices]
    bs = [np.ascontiguousarray(b, dtype=np.float32) for b in biases]
    fs = [np.ascontiguousarray(f, dtype=np.float32) for f in factors]
    q = np.ascontiguousarray(quantiles, dtype=np.float32).reshape(-1, 3)
    c = q.shape[0]
    arr = lambda xs: (ctypes.c_void_p * len(xs))(*[x.ctypes.data for x in xs])
    pm, pb, pf = arr(ms), arr(bs), arr(fs)
    stride = ctypes.c_int32(0)
    check(_L().pcc_factorized_cdf_tables(pm, pb, pf, _np_ptr(q), c, None, ctypes.byref(stride), None, None),
          "pcc_factorized_cdf_tables")
    cdf = np.zeros((c, stride.value), dtype=np.int32)
    length = np.zeros(c, dtype=np.int32)
    offset = np.zeros(c, dtype=np.int32)
    check(_L().pcc_factorized_cdf_tables(pm, pb, pf, _np_ptr(q), c, _np_ptr(cdf), ctypes.byref(stride),
                                         _np_ptr(length), _np_ptr(offset)), "pcc_factorized_cdf_tables")
    return cdf, length, offset
