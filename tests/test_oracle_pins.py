"""Pins of the oracle itself (the reference holds no golden vectors):
  * sparse convolution == torch.nn.functional.conv3d / conv_transpose3d on the zero-filled grid
    (Appendix A3-A5: offset order, stride-2 floor rule, transposed weight direction)
  * rANS: hand-computed known answers + pure-Python big-integer restatement
  * Gaussian tables against scipy.stats.norm
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import random_cloud


def _dense(ot, size, ts=1):
    g = torch.zeros((1, ot.F.shape[1], size, size, size), dtype=torch.float64)
    c = ot.C[:, 1:] // ts
    g[0, :, c[:, 2], c[:, 1], c[:, 0]] = torch.from_numpy(ot.F.T.astype(np.float64))  # [C, z, y, x]
    return g


def _weight_dense(W, ks):
    K, cin, cout = W.shape  # k = dx' + ks*dy' + ks^2*dz'  ->  w[co, ci, dz', dy', dx']
    return torch.from_numpy(W.astype(np.float64)).reshape(ks, ks, ks, cin, cout).permute(4, 3, 0, 1, 2).contiguous()


def _sample(dense, coords, ts):
    c = coords[:, 1:] // ts
    return dense[0][:, c[:, 2], c[:, 1], c[:, 0]].T.numpy()


def test_conv_stride1_equals_dense_conv3d(oracle):
    rng = np.random.default_rng(0)
    c = random_cloud(400, 10, 1, 1, 1)
    x = oracle.prims.OT(c, rng.standard_normal((c.shape[0], 5)).astype(np.float32))
    W = rng.standard_normal((27, 5, 7)).astype(np.float32)
    b = rng.standard_normal((1, 7)).astype(np.float32)
    out = oracle.prims.conv(x, W, b, 3, 1)
    dense = F.conv3d(_dense(x, 12), _weight_dense(W, 3), torch.from_numpy(b[0].astype(np.float64)), padding=1)
    assert np.allclose(out.F, _sample(dense, out.C, 1), atol=1e-4)
    assert np.array_equal(out.C, x.C)


def test_conv_stride2_equals_dense_conv3d(oracle):
    rng = np.random.default_rng(1)
    c = random_cloud(500, 12, 2, 1, 2)  # coordinates in [2, 14)
    x = oracle.prims.OT(c, rng.standard_normal((c.shape[0], 4)).astype(np.float32))
    W = rng.standard_normal((27, 4, 6)).astype(np.float32)
    out = oracle.prims.conv(x, W, None, 3, 2)
    assert out.ts == 2 and (out.C[:, 1:] % 2 == 0).all()
    assert np.array_equal(np.unique(out.C, axis=0), np.unique(np.concatenate([c[:, :1], c[:, 1:] // 2 * 2], 1), axis=0))
    # parent p (even) sees fine voxels p-1, p, p+1: dense conv with padding 1, sampled at even sites
    dense = F.conv3d(_dense(x, 16), _weight_dense(W, 3), None, padding=1)
    assert np.allclose(out.F, _sample(dense, out.C, 1), atol=1e-4)


def test_stride_floor_rule_for_negative_coordinates(oracle):
    c = np.array([[0, -1, -2, -3], [0, -4, 0, 1], [0, 3, 2, -1]], dtype=np.int32)
    s = oracle.prims.stride_coords(c, 2)
    want = np.unique(np.array([[0, -2, -2, -4], [0, -4, 0, 0], [0, 2, 2, -2]], dtype=np.int32), axis=0)
    assert np.array_equal(s, want)


@pytest.mark.parametrize("ks", [2, 3])
def test_generative_transpose_equals_dense_conv_transpose3d(oracle, ks):
    rng = np.random.default_rng(ks)
    c = random_cloud(60, 5, 3, 1, 1)
    c[:, 1:] *= 2  # tensor stride 2
    x = oracle.prims.OT(c, rng.standard_normal((c.shape[0], 3)).astype(np.float32), 2)
    W = rng.standard_normal((ks ** 3, 3, 4)).astype(np.float32)
    out = oracle.prims.conv_transpose(x, W, None, ks)
    assert out.ts == 1
    # out[q] += in[p] @ W[k] for q = p + d_k: conv_transpose3d with stride 2 places input p at 2p'
    # and adds w[ci, co, kz, ky, kx] at offset (kz, ky, kx) - padding
    wt = torch.from_numpy(W.astype(np.float64)).reshape(ks, ks, ks, 3, 4).permute(3, 4, 0, 1, 2).contiguous()
    pad = 1 if ks == 3 else 0
    dense = F.conv_transpose3d(_dense(x, 8, ts=2), wt, None, stride=2, padding=pad, output_padding=1 if ks == 3 else 0)
    got = _sample(dense, out.C, 1)
    assert np.allclose(out.F, got, atol=1e-4)
    # the generated set is exactly the support of the dense result's reachable sites
    assert out.n == np.unique(np.concatenate([(c[:, None, 1:] + oracle.prims.kernel_offsets(ks)[None]).reshape(-1, 3)], 0), axis=0).shape[0]
    if ks == 2:
        assert out.n == 8 * c.shape[0]  # disjoint children (Appendix A5)


def test_features_at_is_gather_or_zero(oracle):
    c = random_cloud(50, 6, 4, 1, 0)
    x = oracle.prims.OT(c, np.arange(c.shape[0] * 2, dtype=np.float32).reshape(-1, 2))
    q = np.concatenate([x.C[:5], np.array([[0, 100, 100, 100]], np.int32)])
    got = oracle.prims.features_at(x, q.astype(np.float32))
    assert np.array_equal(got[:5], x.F[:5]) and (got[5] == 0).all()


def test_rans_known_answers(oracle):
    """Hand-computed with the algorithm of SURVEY.md Appendix B3."""
    E = oracle.entropy
    cdf = np.array([[0, 16384, 49152, 65536, 0]], dtype=np.int32)  # 3 slots; slot 2 is the escape
    length, offset = np.array([4], np.int32), np.array([0], np.int32)
    L = 1 << 31
    # empty message: the flushed initial state, low word first
    assert E.rans_encode(np.zeros(0, np.int32), np.zeros(0, np.int32), cdf, length, offset) == \
        np.array([L & 0xFFFFFFFF, L >> 32], dtype="<u4").tobytes()
    # one symbol v=1 (start 16384, freq 32768): x = ((L // f) << 16) + (L % f) + start
    x = ((L // 32768) << 16) + (L % 32768) + 16384
    assert E.rans_encode(np.array([1], np.int32), np.array([0], np.int32), cdf, length, offset) == \
        np.array([x & 0xFFFFFFFF, x >> 32], dtype="<u4").tobytes()
    # escape: symbol 5 -> v >= max_value(2): raw = 2*(5-2) = 6, one nibble; pushes [slot 2][count 1][nibble 6]
    x = L
    for val in (6, 1):  # reversed push order: nibble, then count
        x = (x << 4) | val
    x = ((x // 16384) << 16) + (x % 16384) + 49152
    words = [x & 0xFFFFFFFF, x >> 32]
    got = E.rans_encode(np.array([5], np.int32), np.array([0], np.int32), cdf, length, offset)
    assert got == np.array(words, dtype="<u4").tobytes()
    assert E.rans_decode(got, np.array([0], np.int32), cdf, length, offset).tolist() == [5]
    # negative escape: -1 -> raw = 1 -> decodes back
    got = E.rans_encode(np.array([-1, 0, 7], np.int32), np.zeros(3, np.int32), cdf, length, offset)
    assert E.rans_decode(got, np.zeros(3, np.int32), cdf, length, offset).tolist() == [-1, 0, 7]


def test_rans_c_equals_python_bigint(oracle):
    E = oracle.entropy
    gc = E.GaussianConditionalOracle()
    rng = np.random.default_rng(3)
    for spread in (1.0, 30.0, 5000.0):
        sym = np.rint(rng.standard_normal(400) * spread).astype(np.int32)
        idx = rng.integers(0, 64, size=400).astype(np.int32)
        assert E.rans_encode(sym, idx, gc.quantized_cdf, gc.cdf_length, gc.offset) == \
            E.rans_encode_py(sym, idx, gc.quantized_cdf, gc.cdf_length, gc.offset)


def test_gaussian_tables_against_scipy(oracle):
    from scipy.stats import norm
    gc = oracle.entropy.GaussianConditionalOracle()
    assert gc.scale_table.shape == (64,) and abs(gc.scale_table[0] - 0.11) < 1e-6 and abs(gc.scale_table[-1] - 256) < 1e-3
    for i in (0, 20, 63):
        s = float(gc.scale_table[i])
        center = -int(gc.offset[i])
        assert center == int(np.ceil(np.float32(s) * np.float32(-norm.ppf(1e-9 / 2))))
        n = int(gc.cdf_length[i]) - 2
        ks = np.arange(n) - center
        pmf = norm.cdf((ks + 0.5) / s) - norm.cdf((ks - 0.5) / s)
        freq = np.diff(gc.quantized_cdf[i, :n + 2])[:n] / 65536.0
        assert np.abs(freq - pmf).max() < 2e-4  # quantisation to 16 bits + frequency stealing
    # build_indexes: boundaries are inclusive (s <= table[j])
    t = gc.scale_table
    idx = gc.build_indexes(np.array([0.0, -5.0, t[0], t[1], np.nextafter(t[1], np.float32(9)), 300.0], np.float32))
    assert idx.tolist() == [0, 0, 0, 1, 2, 63]


def test_topk_tie_break_convention(oracle):
    v = np.array([1.0, 3.0, 3.0, 2.0, 3.0, -0.0, 0.0], np.float32)
    assert np.nonzero(oracle.prims.topk_mask(v, 2))[0].tolist() == [1, 2]
    assert np.nonzero(oracle.prims.topk_mask(v, 5))[0].tolist() == [0, 1, 2, 3, 4]
    assert np.nonzero(oracle.prims.topk_mask(v, 6))[0].tolist() == [0, 1, 2, 3, 4, 6]  # +0 outranks -0


def test_scale_table_is_compressai_formula(pcc, oracle):
    """compressai.models.utils.get_scale_table is torch.exp(torch.linspace(log .11, log 256, 64)) in float32;
    the oracle and the product must both use exactly that (a float64 evaluation differs in the last bit of
    most entries, which moved one CDF index on the 799,458-point ablation frame)."""
    import math
    import torch
    want = torch.exp(torch.linspace(math.log(0.11), math.log(256), 64))
    assert np.array_equal(oracle.entropy.get_scale_table().view(np.uint32), want.numpy().view(np.uint32))
    assert torch.equal(pcc.compressai_compat.get_scale_table(), want)
    m = pcc.ColorModel(pcc.OURS_CONFIG)
    m.update()
    assert torch.equal(m.entropy_model.gaussian_conditional.scale_table, want)
    assert np.array_equal(oracle.entropy.GaussianConditionalOracle().quantized_cdf,
                          m.entropy_model.gaussian_conditional._quantized_cdf.numpy())
