"""ORACLE — TEST INFRASTRUCTURE ONLY.  **parity unpinned** (no golden vectors exist upstream).

Sparse-tensor primitives of the reference's encode/decode path, restated on numpy arrays over
the C primitives in ``pcc_oracle.cpp``.  Semantics: SURVEY.md Appendix A (MinkowskiEngine 0.5.x
behaviour, [upstream-recalled]); each function cites the reference call sites it serves.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this package; the product never does.

Row-order convention: every tensor the oracle creates is in canonical order — lexicographic
(b, x, y, z), the order of ``utils.sort_tensor`` / ``sort_points`` (/root/reference/utils.py:155-204).
ME's own row order is unspecified, and nothing observable on the path depends
on it except top-k tie-breaking, for which the convention is "larger logit, then smaller canonical
row index".
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    """Compile the C primitives (g++ via oracle/Makefile)."""
    src = os.path.join(_HERE, "pcc_oracle.cpp")
    stale = (not os.path.exists(_LIB_PATH)) or (
        os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(_LIB_PATH))
    if force or stale:
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        build()
    L = ctypes.CDLL(_LIB_PATH)
    P = ctypes.c_void_p
    i64, i32 = ctypes.c_int64, ctypes.c_int
    L.orc_num_threads.restype = i32
    L.orc_set_num_threads.argtypes = [i32]
    L.orc_pack_keys.argtypes = [P, i64, P]
    L.orc_unpack_keys.argtypes = [P, i64, P]
    L.orc_kernel_map.argtypes = [P, i64, P, i64, P, i32, P]
    L.orc_lookup.argtypes = [P, i64, P, i64, P]
    L.orc_conv.argtypes = [P, i64, i32, P, i64, i32, P, i32, P, P]
    L.orc_conv_scalar.argtypes = [P, i64, i32, P, i64, i32, P, i32, P, P]
    L.orc_pmf_to_quantized_cdf.argtypes = [P, i32, i32, P]
    L.orc_pmf_to_quantized_cdf.restype = i32
    L.orc_rans_encode.argtypes = [P, P, i64, P, i32, P, P, i32, P, i64]
    L.orc_rans_encode.restype = i64
    L.orc_rans_decode.argtypes = [P, i64, P, i64, P, i32, P, P, i32, P]
    L.orc_rans_decode.restype = i32
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    lib().orc_set_num_threads(int(n))


# ------------------------------------------------------------------ coordinates
COORD_MIN, COORD_MAX = -32768, 32767


def pack_keys(coords):
    """[N,4] int32 (b,x,y,z) -> uint64 keys whose unsigned order is lexicographic (b,x,y,z)."""
    coords = np.ascontiguousarray(coords, dtype=np.int32)
    assert coords.ndim == 2 and coords.shape[1] == 4
    if coords.size:
        assert coords[:, 1:].min() >= COORD_MIN and coords[:, 1:].max() <= COORD_MAX
        assert coords[:, 0].min() >= 0 and coords[:, 0].max() < 65536
    keys = np.empty(coords.shape[0], dtype=np.uint64)
    lib().orc_pack_keys(_p(coords), coords.shape[0], _p(keys))
    return keys


def unpack_keys(keys):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    coords = np.empty((keys.shape[0], 4), dtype=np.int32)
    lib().orc_unpack_keys(_p(keys), keys.shape[0], _p(coords))
    return coords


def kernel_offsets(kernel_size):
    """Kernel offsets in ME's kernel-index order: first spatial axis
    fastest; odd sizes centred ({-1,0,1}), even sizes non-negative ({0,1})."""
    if kernel_size == 1:
        r = [0]
    elif kernel_size == 2:
        r = [0, 1]
    elif kernel_size == 3:
        r = [-1, 0, 1]
    else:
        raise ValueError(kernel_size)
    return np.array([(dx, dy, dz) for dz in r for dy in r for dx in r], dtype=np.int32)


class OT:
    """Oracle sparse tensor: canonical-ordered unique coordinates + features.

    Restates ME.SparseTensor (72 call sites, e.g. model/model.py:122,188):
    coordinates [N,4] int32 batch-first, float coordinates floored, tensor_stride recorded.
    Duplicate coordinates are rejected (the codec never passes any).
    """

    def __init__(self, coords, feats, ts=1, _sorted=False):
        coords = np.asarray(coords)
        if coords.dtype.kind == "f":
            coords = np.floor(coords)  # model/model.py:184-188 passes float coordinates
        coords = np.ascontiguousarray(coords, dtype=np.int32)
        feats = np.ascontiguousarray(feats, dtype=np.float32)
        assert feats.shape[0] == coords.shape[0]
        keys = pack_keys(coords)
        if not _sorted:
            perm = np.argsort(keys, kind="stable")
            keys, coords, feats = keys[perm], coords[perm], feats[perm]
        assert keys.size < 2 or (keys[1:] > keys[:-1]).all(), "duplicate or unsorted coordinates"
        self.C, self.F, self.keys = coords, np.ascontiguousarray(feats), np.ascontiguousarray(keys)
        self.ts = int(ts[0]) if isinstance(ts, (list, tuple)) else int(ts)

    @property
    def n(self):
        return self.C.shape[0]

    def with_feats(self, feats):
        return OT(self.C, feats, self.ts, _sorted=True)


def lookup(x_keys, q_coords):
    """rows of query coordinates in a canonical key set, -1 when absent."""
    q_keys = pack_keys(q_coords)
    rows = np.empty(q_keys.shape[0], dtype=np.int32)
    lib().orc_lookup(_p(x_keys), x_keys.shape[0], _p(q_keys), q_keys.shape[0], _p(rows))
    return rows


def features_at(x, q_coords):
    """SparseTensor.features_at_coordinates for on-grid integer queries == gather-or-zero
    (model/blocks.py:37,50, model/transforms.py:96,124,262,
    model/entropy_models.py:364,401)."""
    q = np.asarray(q_coords)
    if q.dtype.kind == "f":
        assert (q == np.floor(q)).all(), "oracle restates on-grid queries only"
    rows = lookup(x.keys, q.astype(np.int32))
    out = np.zeros((rows.shape[0], x.F.shape[1]), dtype=np.float32)
    hit = rows >= 0
    out[hit] = x.F[rows[hit]]
    return out


def stride_coords(coords, ts_out):
    """Output coordinates of a stride-2 convolution: unique(floor(c / ts') * ts') per spatial axis
    (SURVEY.md Appendix A4), canonical order."""
    c = np.asarray(coords, dtype=np.int64).copy()
    c[:, 1:] = np.floor_divide(c[:, 1:], ts_out) * ts_out
    keys = np.unique(pack_keys(c.astype(np.int32)))
    return unpack_keys(keys)


def expand_coords(coords, kernel_size, ts_out):
    """Output coordinates of a generative transposed convolution:
    unique{ p + d_k * ts' }, canonical order."""
    offs = kernel_offsets(kernel_size).astype(np.int64) * ts_out
    c = np.asarray(coords, dtype=np.int64)
    cand = np.repeat(c[:, None, :], offs.shape[0], axis=1)
    cand[:, :, 1:] += offs[None, :, :]
    keys = np.unique(pack_keys(cand.reshape(-1, 4).astype(np.int32)))
    return unpack_keys(keys)


def kernel_map(in_keys, out_coords, offs):
    """nbr[k, o] = row of (out_coords[o] + offs[k]) in the canonical input set, or -1."""
    out_coords = np.ascontiguousarray(out_coords, dtype=np.int32)
    offs = np.ascontiguousarray(offs, dtype=np.int32)
    K, n_out = offs.shape[0], out_coords.shape[0]
    nbr = np.empty((K, n_out), dtype=np.int32)
    lib().orc_kernel_map(_p(in_keys), in_keys.shape[0], _p(out_coords), n_out, _p(offs), K, _p(nbr))
    return nbr


def kernel_map_pairs(nbr):
    """ME-style kernel map: per kernel index k the (in_row, out_row) pairs, out_row ascending."""
    pairs = []
    for k in range(nbr.shape[0]):
        o = np.nonzero(nbr[k] >= 0)[0].astype(np.int32)
        pairs.append(np.stack([nbr[k][o], o], axis=1))
    return pairs


def conv_apply(in_feats, nbr, W, bias, scalar=False):
    """out[o] = sum_k in[nbr[k,o]] @ W[k] (+ bias) in the exact fp32 order of pcc_oracle.cpp."""
    in_feats = np.ascontiguousarray(in_feats, dtype=np.float32)
    W = np.ascontiguousarray(W, dtype=np.float32)
    if W.ndim == 2:  # ME stores kernel_size=1 weights as [Cin, Cout]
        W = W[None]
    K, cin, cout = W.shape
    assert nbr.shape[0] == K and in_feats.shape[1] == cin, (nbr.shape, W.shape, in_feats.shape)
    n_out = nbr.shape[1]
    b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32).reshape(-1)
    out = np.empty((n_out, cout), dtype=np.float32)
    fn = lib().orc_conv_scalar if scalar else lib().orc_conv
    fn(_p(in_feats), in_feats.shape[0], cin, _p(np.ascontiguousarray(nbr)), n_out, K, _p(W), cout,
       _p(b), _p(out))
    return out


def conv(x, W, bias, kernel_size, stride=1):
    """ME.MinkowskiConvolution (57 call sites)."""
    offs = kernel_offsets(kernel_size) * x.ts
    if stride == 1:
        out_coords, ts_out = x.C, x.ts
    else:
        assert stride == 2
        ts_out = x.ts * 2
        out_coords = stride_coords(x.C, ts_out)
    nbr = kernel_map(x.keys, out_coords, offs)
    return OT(out_coords, conv_apply(x.F, nbr, W, bias), ts_out, _sorted=True)


def conv_transpose(x, W, bias, kernel_size, stride=2, out_coords=None):
    """ME.MinkowskiGenerativeConvolutionTranspose (model/blocks.py:84, model/entropy_models.py:286,290)
    and ME.MinkowskiConvolutionTranspose on a fresh coordinate manager (model/entropy_models.py:298,302),
    which generates its output map the same way.  out[q] += in[p] @ W[k] for
    q = p + d_k * ts'.  `out_coords` restricts evaluation to a given target set (an existing map)."""
    assert stride == 2 and x.ts % 2 == 0
    ts_out = x.ts // 2
    if out_coords is None:
        out_coords = expand_coords(x.C, kernel_size, ts_out)
    offs = -kernel_offsets(kernel_size) * ts_out  # p = q - d_k * ts'
    nbr = kernel_map(x.keys, out_coords, offs)
    return OT(out_coords, conv_apply(x.F, nbr, W, bias), ts_out, _sorted=True)


def relu(a):
    return np.maximum(a, np.float32(0))


def leaky_relu(a, slope=0.01):
    """ME.MinkowskiLeakyReLU default negative slope 0.01."""
    a = np.asarray(a, dtype=np.float32)
    return np.where(a > 0, a, a * np.float32(slope)).astype(np.float32)


def float_order_key(v):
    """Monotone uint32 image of float32 values (total order; -0 < +0)."""
    u = np.ascontiguousarray(v, dtype=np.float32).view(np.uint32)
    return np.where(u & np.uint32(0x80000000), ~u, u | np.uint32(0x80000000)).astype(np.uint32)


def topk_mask(logits, k):
    """GenerativeUpBlock._topk_prediction for one batch element (model/blocks.py:130-150):
    boolean mask of the k largest logits; ties -> smaller canonical row index (Appendix A11)."""
    n = logits.shape[0]
    k = int(k)
    assert 0 <= k <= n, "torch.topk raises when k exceeds the candidate count"
    key = float_order_key(logits).astype(np.int64)
    order = np.argsort(-key, kind="stable")  # descending key, ascending row among equals
    mask = np.zeros(n, dtype=bool)
    mask[order[:k]] = True
    return mask
