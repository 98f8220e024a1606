"""GPU parity of the integer/byte primitives against the CPU oracle (bit-exact), through the C-ABI."""
import numpy as np
import pytest
import torch

from conftest import random_cloud

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _keys_np(t):
    return t.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("n,extent,lo,batch", [(1, 4, 0, 1), (300, 10, -5, 1), (5000, 40, 0, 2), (200000, 300, -20, 1)])
def test_pack_sort_unpack(pcc, oracle, n, extent, lo, batch):
    c = random_cloud(n, extent, 1, batch, lo)
    ct = torch.from_numpy(c).to(DEV)
    keys, status = pcc.ops.pack_coords(ct)
    assert int(status.item()) == 0
    ref = oracle.prims.pack_keys(c)
    assert np.array_equal(_keys_np(keys), ref)
    skeys, perm = pcc.ops.sort_keys(keys)
    order = np.argsort(ref, kind="stable")
    assert np.array_equal(_keys_np(skeys), ref[order])
    assert np.array_equal(perm.cpu().numpy(), order.astype(np.int32))
    assert np.array_equal(pcc.ops.unpack_keys(skeys).cpu().numpy(), c[order])


def test_pack_out_of_range_and_duplicates(pcc):
    bad = torch.tensor([[0, 40000, 0, 0]], dtype=torch.int32, device=DEV)
    with pytest.raises(pcc.PccError):
        pcc.ops.canonical_map(bad)
    dup = torch.tensor([[0, 1, 2, 3], [0, 1, 2, 3]], dtype=torch.int32, device=DEV)
    with pytest.raises(pcc.PccError):
        pcc.ops.canonical_map(dup)


def test_empty_inputs(pcc):
    e = torch.zeros((0, 4), dtype=torch.int32, device=DEV)
    m, perm = pcc.ops.canonical_map(e)
    assert m.n == 0 and perm.numel() == 0
    assert m.stride2().n == 0
    rows = m.lookup(torch.zeros(0, dtype=torch.int64, device=DEV))
    assert rows.numel() == 0


@pytest.mark.parametrize("n,extent,lo", [(50, 6, -3), (4000, 30, 0), (150000, 200, -10)])
def test_hash_lookup(pcc, oracle, n, extent, lo):
    c = random_cloud(n, extent, 2, 1, lo)
    q = random_cloud(n, extent + 4, 3, 1, lo - 2)
    m = pcc.ops.make_map(torch.from_numpy(c).to(DEV))
    qk, _ = pcc.ops.pack_coords(torch.from_numpy(q).to(DEV))
    rows = m.lookup(qk).cpu().numpy()
    ck = oracle.prims.pack_keys(c)
    order = np.argsort(ck)
    ref_sorted = oracle.prims.lookup(ck[order], q)
    ref = np.where(ref_sorted >= 0, order[np.maximum(ref_sorted, 0)], -1)
    assert np.array_equal(rows, ref.astype(np.int32))


@pytest.mark.parametrize("ts", [1, 2, 8])
@pytest.mark.parametrize("n,extent,lo", [(40, 5, -4), (6000, 40, -7), (200000, 250, 0)])
def test_stride_map(pcc, oracle, ts, n, extent, lo):
    c = random_cloud(n, extent, 4, 2, lo)
    c[:, 1:] *= ts
    m, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), ts)
    s = m.stride2()
    ref = oracle.prims.stride_coords(c, 2 * ts)
    assert s.ts == 2 * ts
    assert np.array_equal(s.coords.cpu().numpy(), ref)


@pytest.mark.parametrize("ks", [2, 3])
@pytest.mark.parametrize("n,extent,lo,ts", [(1, 2, 0, 2), (30, 4, -2, 2), (5000, 30, -5, 4), (60000, 90, 0, 16)])
def test_expand_map(pcc, oracle, ks, n, extent, lo, ts):
    c = random_cloud(n, extent, 5, 2, lo)
    c[:, 1:] *= ts
    m, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), ts)
    e = m.expand(ks)
    ref = oracle.prims.expand_coords(c, ks, ts // 2)
    assert e.ts == ts // 2
    assert np.array_equal(e.coords.cpu().numpy(), ref)


def _oracle_sorted(oracle, c, ts):
    keys = np.sort(oracle.prims.pack_keys(c))
    return keys, oracle.prims.unpack_keys(keys)


@pytest.mark.parametrize("ks,groups", [(1, None), (2, None), (3, 1), (3, 3), (3, 5)])
@pytest.mark.parametrize("n,extent,ts", [(60, 6, 1), (7000, 35, 2), (150000, 150, 1)])
def test_kernel_map_conv(pcc, oracle, ks, groups, n, extent, ts):
    pcc.ops.set_tile_segment_rows(1 << 16 if n > 100000 else None)  # multi-segment case
    c = random_cloud(n, extent, 6, 1, -3)
    c[:, 1:] *= ts
    m, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), ts)
    keys, cs = _oracle_sorted(oracle, c, ts)
    offs = oracle.prims.kernel_offsets(ks) * ts
    # stride 1
    km = m.kernel_map(m, ks, groups=groups)
    ref = oracle.prims.kernel_map(keys, cs, offs)
    assert np.array_equal(km.natural_nbr().cpu().numpy(), ref)
    # tile rows: per offset group, `order` is a permutation (sorted by the group's bits of the neighbour
    # mask inside lexicographic segments) and the tap mask is the OR over the rows of each 128-row tile
    n_out = ref.shape[1]
    tiles = (n_out + 127) // 128
    tm = km.tapmask.cpu().numpy().view(np.uint32)
    nbr_t = km.nbr[:, :km.n_out].cpu().numpy()
    assert km.n_groups == (1 if (groups is None or km.order is None) else groups)
    for g, (lo, hi) in enumerate(km.group_ranges()):
        order = np.arange(n_out) if km.order is None else km.order.view(km.n_groups, -1)[g].cpu().numpy()
        assert np.array_equal(np.sort(order), np.arange(n_out))
        tiled = ref[lo:hi][:, order]
        assert np.array_equal(nbr_t[lo:hi], tiled)
        if km.order is not None:
            seg = pcc.ops.tile_segment_rows()
            rowmask = ((ref[lo:hi] >= 0).astype(np.int64) << np.arange(hi - lo)[:, None]).sum(0)
            sortkey = ((np.arange(n_out) // seg) << (hi - lo)) | rowmask
            assert np.array_equal(order, np.argsort(sortkey, kind="stable"))
        for t in range(tiles):
            want = 0
            for k in range(lo, hi):
                if (tiled[k - lo, t * 128:(t + 1) * 128] >= 0).any():
                    want |= 1 << k
            assert tm[g * tiles + t] == want
        if km.tile_order is not None:  # launch order: a permutation of the tiles, most active offsets first
            to = km.tile_order.view(km.n_groups, -1)[g].cpu().numpy()
            assert np.array_equal(np.sort(to), np.arange(tiles))
            pc = np.array([bin(int(v)).count("1") for v in tm[g * tiles:(g + 1) * tiles]])[to]
            assert np.all(pc[1:] <= pc[:-1])
    # stride 2
    s = m.stride2()
    km2 = m.kernel_map(s, ks)
    ref2 = oracle.prims.kernel_map(keys, oracle.prims.stride_coords(c, 2 * ts), offs)
    assert np.array_equal(km2.natural_nbr().cpu().numpy(), ref2)
    pcc.ops.set_tile_segment_rows(None)


@pytest.mark.parametrize("ks", [2, 3])
@pytest.mark.parametrize("n,extent,ts", [(50, 5, 2), (20000, 60, 8)])
def test_kernel_map_transpose(pcc, oracle, ks, n, extent, ts):
    c = random_cloud(n, extent, 7, 1, -2)
    c[:, 1:] *= ts
    m, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), ts)
    keys, cs = _oracle_sorted(oracle, c, ts)
    e = m.expand(ks)
    km = m.kernel_map(e, ks, transpose=True)
    ref = oracle.prims.kernel_map(keys, oracle.prims.expand_coords(c, ks, ts // 2),
                                  -oracle.prims.kernel_offsets(ks) * (ts // 2))
    got = km.natural_nbr().cpu().numpy()
    assert np.array_equal(got, ref)
    assert (got >= 0).sum() == ks ** 3 * m.n  # generative: pairs = K * N_in exactly (Appendix A5)


@pytest.mark.parametrize("m,k", [(1, 1), (10, 0), (10, 10), (1000, 37), (100000, 25000), (1 << 20, 300000)])
def test_topk_select(pcc, oracle, m, k):
    rng = np.random.default_rng(m + k)
    v = rng.standard_normal(m).astype(np.float32)
    v[rng.integers(0, m, size=m // 3)] = np.float32(0.25)  # heavy ties
    if m > 5:
        v[:3] = [0.0, -0.0, 0.25]
    logits = torch.from_numpy(np.stack([v, -v], 1).copy()).to(DEV)
    sel = pcc.ops.topk_rows(logits, k).cpu().numpy()
    ref = np.nonzero(oracle.prims.topk_mask(v, k))[0].astype(np.int32)
    assert np.array_equal(sel, ref)


def test_topk_rejects_k_above_m(pcc):
    with pytest.raises(pcc.PccError):
        pcc.ops.topk_rows(torch.zeros((4, 1), device=DEV), 5)
