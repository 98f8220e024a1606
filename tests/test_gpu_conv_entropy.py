"""GPU parity of the exact-fp32 sparse convolution (bit-exact vs the oracle's fmaf order) and of
the entropy-stage kernels, through the C-ABI."""
import numpy as np
import pytest
import torch

from conftest import random_cloud

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(pcc, oracle, n, extent, ts, seed):
    c = random_cloud(n, extent, seed, 1, -2)
    c[:, 1:] *= ts
    m, perm = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), ts)
    keys = np.sort(oracle.prims.pack_keys(c))
    return m, keys, oracle.prims.unpack_keys(keys)


@pytest.mark.parametrize("cin,cout", [(1, 1), (2, 2), (4, 2), (4, 64), (2, 16), (16, 2), (32, 3), (64, 32),
                                      (64, 128), (128, 128), (128, 256), (192, 256), (20, 35), (128, 4)])
def test_conv_exact_bitwise(pcc, oracle, cin, cout):
    rng = np.random.default_rng(cin * 1000 + cout)
    m, keys, cs = _setup(pcc, oracle, 3000, 22, 1, cin + cout)
    n = m.n
    x = rng.standard_normal((n, cin)).astype(np.float32)
    W = (rng.standard_normal((27, cin, cout)) / np.sqrt(cin * 9)).astype(np.float32)
    b = rng.standard_normal((1, cout)).astype(np.float32)
    nbr = oracle.prims.kernel_map(keys, cs, oracle.prims.kernel_offsets(3))
    ref = oracle.prims.conv_apply(x, nbr, W, b)
    km = m.kernel_map(m, 3)
    got = pcc.ops.conv(torch.from_numpy(x).to(DEV), km, torch.from_numpy(W).to(DEV), torch.from_numpy(b).to(DEV),
                       mode=pcc.ops.MODE_EXACT)
    assert np.array_equal(got.cpu().numpy().view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("ks,stride,transpose", [(1, 1, False), (3, 2, False), (3, 2, True), (2, 2, True)])
def test_conv_variants_bitwise(pcc, oracle, ks, stride, transpose):
    rng = np.random.default_rng(ks * 10 + stride)
    ts = 4
    m, keys, cs = _setup(pcc, oracle, 5000, 26, ts, 11)
    cin, cout = 32, 64
    x = rng.standard_normal((m.n, cin)).astype(np.float32)
    W = (rng.standard_normal((ks ** 3, cin, cout)) / 8).astype(np.float32)
    xo = oracle.prims.OT(cs, x, ts, _sorted=True)
    if transpose:
        ref = oracle.prims.conv_transpose(xo, W, None, ks)
        out_map = m.expand(ks)
    else:
        ref = oracle.prims.conv(xo, W, None, ks, stride)
        out_map = m.stride2() if stride == 2 else m
    km = m.kernel_map(out_map, ks, transpose=transpose)
    Wt = torch.from_numpy(W if ks > 1 else W[0]).to(DEV)
    got = pcc.ops.conv(torch.from_numpy(x).to(DEV), km, Wt, None)
    assert np.array_equal(out_map.coords.cpu().numpy(), ref.C)
    assert np.array_equal(got.cpu().numpy().view(np.uint32), ref.F.view(np.uint32))


@pytest.mark.parametrize("cin,cout,film_c", [(24, 48, 0), (24, 48, 1), (24, 48, 48),
                                             (64, 4, 0), (64, 4, 1), (64, 4, 4), (128, 3, 3), (40, 2, 1), (16, 1, 1),
                                             # conv_exact_tma_kernel <64,4> / <128,8>: 16-byte epilogue (conv_epilogue4)
                                             (64, 64, 64), (32, 64, 1), (64, 128, 128), (128, 256, 0), (64, 36, 36)])
@pytest.mark.parametrize("act", [0, 1, 2])
def test_conv_epilogue_bitwise(pcc, oracle, cin, cout, film_c, act):
    """bias -> activation -> FiLM (x*beta+gamma, model/blocks.py:40) -> residual (model/blocks.py:50);
    cout <= 4 with cin % 4 == 0 runs the narrow-output kernel (conv_head_kernel); cin % 32 == 0 with a 64- or
    128-wide tile the TMA kernel, whose epilogue moves four columns per access."""
    rng = np.random.default_rng(film_c + act + cin)
    m, keys, cs = _setup(pcc, oracle, 2500, 20, 1, 5)
    n = m.n
    x = rng.standard_normal((n, cin)).astype(np.float32)
    W = (rng.standard_normal((27, cin, cout)) / 10).astype(np.float32)
    b = rng.standard_normal((1, cout)).astype(np.float32)
    nbr = oracle.prims.kernel_map(keys, cs, oracle.prims.kernel_offsets(3))
    ref = oracle.prims.conv_apply(x, nbr, W, b)
    if act == 1:
        ref = oracle.prims.relu(ref)
    elif act == 2:
        ref = oracle.prims.leaky_relu(ref)
    film = frows = None
    if film_c:
        nf = n // 2
        film = rng.standard_normal((nf, 2 * film_c)).astype(np.float32)
        frows = rng.integers(-1, nf, size=n).astype(np.int32)
        g = np.where(frows[:, None] >= 0, film[np.maximum(frows, 0)], np.float32(0))
        ref = (ref * g[:, :film_c] + g[:, film_c:]).astype(np.float32)
    res = rng.standard_normal((n, cout)).astype(np.float32)
    rrows = rng.integers(-1, n, size=n).astype(np.int32)
    ref = ref + np.where(rrows[:, None] >= 0, res[np.maximum(rrows, 0)], np.float32(0))
    t = lambda a: None if a is None else torch.from_numpy(a).to(DEV)
    got = pcc.ops.conv(t(x), m.kernel_map(m, 3), t(W), t(b), act=act, film=t(film), film_rows=t(frows),
                       residual=t(res), res_rows=t(rrows))
    assert np.array_equal(got.cpu().numpy().view(np.uint32), ref.astype(np.float32).view(np.uint32))


@pytest.mark.parametrize("cin", [1, 2, 3, 4])
@pytest.mark.parametrize("cout,groups", [(1, 1), (2, 3), (3, 1), (16, 1), (20, 3), (64, 1), (128, 1), (130, 3)])
def test_conv_thin_input_bitwise(pcc, oracle, cin, cout, groups):
    """cin <= 4, 27 offsets: conv_thin_kernel (neighbour rows in registers, weights in shared memory), with the
    fused epilogue and with the offsets cut into chained groups; several tiles per CTA-stride and a ragged tail."""
    rng = np.random.default_rng(cin * 100 + cout)
    m, keys, cs = _setup(pcc, oracle, 4000, 24, 1, cin + 7 * cout)
    n = m.n
    x = rng.standard_normal((n, cin)).astype(np.float32)
    W = rng.standard_normal((27, cin, cout)).astype(np.float32)
    b = rng.standard_normal((1, cout)).astype(np.float32)
    nbr = oracle.prims.kernel_map(keys, cs, oracle.prims.kernel_offsets(3))
    ref = oracle.prims.leaky_relu(oracle.prims.conv_apply(x, nbr, W, b))
    film = rng.standard_normal((n, 2 * cout)).astype(np.float32)
    frows = rng.integers(-1, n, size=n).astype(np.int32)
    g = np.where(frows[:, None] >= 0, film[np.maximum(frows, 0)], np.float32(0))
    ref = (ref * g[:, :cout] + g[:, cout:]).astype(np.float32)
    res = rng.standard_normal((n, cout)).astype(np.float32)
    ref = ref + res
    t = lambda a: torch.from_numpy(a).to(DEV)
    km = m.kernel_map(m, 3, groups=groups)
    assert km.n_groups == groups
    got = pcc.ops.conv(t(x), km, t(W), t(b), act=2, film=t(film), film_rows=t(frows), residual=t(res))
    assert np.array_equal(got.cpu().numpy().view(np.uint32), ref.view(np.uint32))
    plain = pcc.ops.conv(t(x), km, t(W), None)
    assert np.array_equal(plain.cpu().numpy().view(np.uint32),
                          oracle.prims.conv_apply(x, nbr, W, None).view(np.uint32))


def test_conv_zero_rows_and_tiny(pcc, oracle):
    m, keys, cs = _setup(pcc, oracle, 1, 2, 1, 3)
    x = torch.ones((m.n, 8), device=DEV)
    W = torch.ones((27, 8, 8), device=DEV)
    out = pcc.ops.conv(x, m.kernel_map(m, 3), W, None)
    assert out.shape == (m.n, 8) and torch.all(out == 8.0)
    e = pcc.ops.CoordMap(torch.zeros(0, dtype=torch.int64, device=DEV), 1, True)
    out = pcc.ops.conv(torch.zeros((0, 8), device=DEV), e.kernel_map(e, 3), W, None)
    assert out.shape == (0, 8)


def test_gather_rows(pcc):
    rng = np.random.default_rng(0)
    for c in (2, 3, 128):
        f = rng.standard_normal((500, c)).astype(np.float32)
        rows = rng.integers(-1, 500, size=900).astype(np.int32)
        add = rng.standard_normal((900, c)).astype(np.float32)
        ref = np.where(rows[:, None] >= 0, f[np.maximum(rows, 0)], np.float32(0))
        got = pcc.ops.gather_rows(torch.from_numpy(f).to(DEV), torch.from_numpy(rows).to(DEV))
        assert np.array_equal(got.cpu().numpy(), ref)
        got = pcc.ops.gather_rows(torch.from_numpy(f).to(DEV), torch.from_numpy(rows).to(DEV), add=torch.from_numpy(add).to(DEV))
        assert np.array_equal(got.cpu().numpy(), add + ref)


# ------------------------------------------------------------------ entropy stage
@pytest.mark.parametrize("n,c", [(1, 128), (33, 5), (1178, 128), (17957, 128)])
def test_gc_quantize_index_dequantize(pcc, oracle, n, c):
    rng = np.random.default_rng(n + c)
    gc = oracle.entropy.GaussianConditionalOracle()
    y = (rng.standard_normal((n, c)) * 6).astype(np.float32)
    y[0, 0] = 2.5  # half-to-even cases
    npar = n + 7
    params = np.concatenate([np.exp(rng.uniform(-4, 6, size=(npar, c))) * rng.choice([-1, 1], size=(npar, c)),
                             rng.standard_normal((npar, c)) * 3], axis=1).astype(np.float32)
    params[:, c:][0, 0] = 0.0
    params[3, :c] = gc.scale_table[:c] if c <= 64 else np.resize(gc.scale_table, c)  # exact table hits
    rows = rng.integers(-1, npar, size=n).astype(np.int32)
    rows[0] = 0
    g = np.where(rows[:, None] >= 0, params[np.maximum(rows, 0)], np.float32(0))
    scales, means = g[:, :c].T, g[:, c:].T
    ref_idx = gc.build_indexes(scales)
    ref_sym = gc.quantize(y.T, means)
    table = torch.from_numpy(gc.scale_table).to(DEV)
    t = lambda a: torch.from_numpy(a).to(DEV)
    sym, idx = pcc.ops.gc_quantize_index(t(y), t(params), t(rows), table, 0.11)
    assert np.array_equal(idx.cpu().numpy(), ref_idx)
    assert np.array_equal(sym.cpu().numpy(), ref_sym)
    idx2 = pcc.ops.gc_build_indexes(t(params), t(rows), n, c, table, 0.11)
    assert np.array_equal(idx2.cpu().numpy(), ref_idx)
    y_hat = pcc.ops.gc_dequantize(sym, t(params), t(rows))
    assert np.array_equal(y_hat.cpu().numpy(), (ref_sym.astype(np.float32) + means).T)


def test_eb_quantize_dequantize(pcc, oracle):
    rng = np.random.default_rng(1)
    n, c = 1178, 128
    z = (rng.standard_normal((n, c)) * 4).astype(np.float32)
    med = rng.standard_normal(c).astype(np.float32)
    sym, z_hat = pcc.ops.eb_quantize(torch.from_numpy(z).to(DEV), torch.from_numpy(med).to(DEV))
    ref = np.rint(z - med[None, :]).astype(np.int32)
    assert np.array_equal(sym.cpu().numpy(), ref.T)
    assert np.array_equal(z_hat.cpu().numpy(), ref.astype(np.float32) + med[None, :])
    back = pcc.ops.eb_dequantize(sym, torch.from_numpy(med).to(DEV))
    assert np.array_equal(back.cpu().numpy(), z_hat.cpu().numpy())


def test_likelihoods_tolerance(pcc, oracle):
    """Floating-point outputs: tolerance 1e-5 absolute on likelihoods in [1e-9, 1] (fp32 erfc /
    sigmoid on the GPU vs float64 in the oracle)."""
    rng = np.random.default_rng(2)
    n, c = 700, 128
    model = pcc.ColorModel(pcc.OURS_CONFIG)
    pcc.codec.init_synthetic(model, 3)
    model.eval()
    eb = model.entropy_model.entropy_bottleneck
    gcm = model.entropy_model.gaussian_conditional
    model.update()
    model.to(DEV)
    # factorised prior
    z = (rng.standard_normal((n, c)) * 5).astype(np.float32)
    ebo = oracle.entropy.EntropyBottleneckOracle({k: v.detach().cpu().numpy() for k, v in eb.named_parameters()})
    z_hat_ref, lik_ref = ebo.forward(z.T)
    z_hat, lik = eb(torch.from_numpy(z).to(DEV).t().unsqueeze(0))
    assert np.array_equal(z_hat[0].cpu().numpy(), z_hat_ref)
    assert np.abs(lik[0].cpu().numpy() - lik_ref).max() < 1e-5
    # gaussian conditional
    y = (rng.standard_normal((n, c)) * 3).astype(np.float32)
    scales = np.exp(rng.uniform(-3, 3, size=(n, c))).astype(np.float32)
    means = rng.standard_normal((n, c)).astype(np.float32)
    gco = oracle.entropy.GaussianConditionalOracle()
    y_hat_ref, ylik_ref = gco.forward(y.T, scales.T, means.T)
    t = lambda a: torch.from_numpy(a).to(DEV).t().unsqueeze(0)
    y_hat, ylik = gcm(t(y), t(scales), means=t(means))
    assert np.array_equal(y_hat[0].cpu().numpy(), y_hat_ref)
    assert np.abs(ylik[0].cpu().numpy() - ylik_ref).max() < 1e-5


def test_color_finalize(pcc, oracle):
    rng = np.random.default_rng(4)
    c = random_cloud(500, 30, 9, 1, 0)
    m, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV))
    f = rng.uniform(-0.2, 1.2, size=(m.n, 3)).astype(np.float32)
    f[0] = [0.5 / 255, 1.5 / 255, 2.5 / 255]
    got = pcc.ops.color_finalize(m.keys, torch.from_numpy(f).to(DEV)).cpu().numpy()
    ref = np.clip(np.rint(f * np.float32(255)), 0, 255).astype(np.float32) / np.float32(255)
    assert np.array_equal(got[:, 3:], ref)
    assert np.array_equal(got[:, :3], m.coords.cpu().numpy()[:, 1:].astype(np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize("seg,groups", [(0, 1), (128, 1), (1024, 1), (1 << 17, 1), (128, 3), (1 << 17, 3), (1 << 30, 9),
                                        (1 << 30, 27)])
def test_conv_independent_of_tile_order(pcc, oracle, seg, groups):
    """The tile-row permutation(s) of the kernel map (rows grouped by neighbour mask, per offset
    group) must not change a single bit of the result, in either mode, with FiLM / residual row
    lookups in the epilogue; grouped maps run one chained pass per group (exact mode only)."""
    ops = pcc.ops
    rng = np.random.default_rng(5)
    old = ops._tile_seg_rows
    try:
        ops.set_tile_segment_rows(seg)
        m, keys, cs = _setup(pcc, oracle, 6000, 30, 1, 77)
        n, cin, cout = m.n, 64, 128
        x = rng.standard_normal((n, cin)).astype(np.float32)
        W = (rng.standard_normal((27, cin, cout)) / np.sqrt(cin * 9)).astype(np.float32)
        b = rng.standard_normal((1, cout)).astype(np.float32)
        film = rng.standard_normal((n // 2, 2 * cout)).astype(np.float32)
        frows = rng.integers(-1, n // 2, size=n).astype(np.int32)
        res = rng.standard_normal((n, cout)).astype(np.float32)
        nbr = oracle.prims.kernel_map(keys, cs, oracle.prims.kernel_offsets(3))
        ref = oracle.prims.relu(oracle.prims.conv_apply(x, nbr, W, b))
        g = np.where(frows[:, None] >= 0, film[np.maximum(frows, 0)], np.float32(0))
        ref = ((ref * g[:, :cout] + g[:, cout:]).astype(np.float32) + res).astype(np.float32)
        km = m.kernel_map(m, 3, groups=groups)
        assert (km.order is None) == (seg == 0) and km.n_groups == groups
        t = lambda a: torch.from_numpy(a).to(DEV)
        got = ops.conv(t(x), km, t(W), t(b), act=1, film=t(film), film_rows=t(frows), residual=t(res),
                       mode=ops.MODE_EXACT)
        assert np.array_equal(got.cpu().numpy().view(np.uint32), ref.view(np.uint32))
        # narrow-output kernel over the same (grouped) map
        W4 = np.ascontiguousarray(W[:, :, :3])
        ref4 = oracle.prims.conv_apply(x, nbr, W4, np.ascontiguousarray(b[:, :3]))
        got4 = ops.conv(t(x), km, t(W4), t(np.ascontiguousarray(b[:, :3])), mode=ops.MODE_EXACT)
        assert np.array_equal(got4.cpu().numpy().view(np.uint32), ref4.astype(np.float32).view(np.uint32))
        if groups == 1:
            g32 = ops.conv(t(x), km, t(W), t(b), act=1, film=t(film), film_rows=t(frows), residual=t(res),
                           mode=ops.MODE_TF32, round_out=False)
            assert (g32 - got).abs().max().item() <= 2e-2
        else:
            with pytest.raises(pcc.PccError):
                ops.conv(t(x), km, t(W), t(b), mode=ops.MODE_TF32)
    finally:
        ops.set_tile_segment_rows(old)
