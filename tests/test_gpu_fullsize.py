"""BASELINE.json's full sizes (configs[1], configs[2]: one 799,458-point 10-bit frame) checked through
size-independent properties — the oracle needs minutes at this size, so it is not run here:

* coordinate maps: sorted, unique, equal to an independent torch restatement (floor / unique);
* kernel maps: centre tap = identity, tap k of row o <-> tap 26 - k of the neighbour (symmetry);
* convolution: conv(2 x) == 2 conv(x) bit for bit (power-of-two scaling commutes with every rounding);
* top-k: exactly k rows, ascending, min(selected) >= max(rejected);
* codec: decoder recovers the encoder's quantised latent bit for bit, the strings do not depend on the
  run (determinism) nor on the input row order, decoded cloud has exactly N unique points with colours
  on the 8-bit grid; over the quality sweep q in {0, 0.5, 1}^2 of configs[1] (corners + centre).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def frame(pcc):
    x = pcc.synthetic.shell_frame(252.0, seed=0)
    assert x.shape[0] == 799458
    return x


@pytest.fixture(scope="module")
def level0(pcc, frame):
    c = torch.cat([torch.zeros((frame.shape[0], 1)), frame[:, :3]], 1).int().to(DEV)
    m, perm = pcc.ops.canonical_map(c, 1)
    return m


def _unpack(keys):
    x = ((keys >> 32) & 0xFFFF) - 32768
    y = ((keys >> 16) & 0xFFFF) - 32768
    z = (keys & 0xFFFF) - 32768
    return x, y, z


def test_maps_full_size(pcc, level0):
    m = level0
    assert m.n == 799458 and bool((m.keys[1:] > m.keys[:-1]).all())
    prev = m
    for ts in (2, 4, 8):
        child = prev.stride2()
        x, y, z = _unpack(prev.keys)
        want = torch.unique((((x // ts) * ts + 32768) << 32) | (((y // ts) * ts + 32768) << 16) | ((z // ts) * ts + 32768))
        assert torch.equal(child.keys, want), f"stride-{ts} map"
        prev = child
    # generative map of the stride-8 level: unique { p + d * 4 }, d in {-1, 0, 1}^3
    x, y, z = _unpack(prev.keys)
    cand = []
    for dx in (-4, 0, 4):
        for dy in (-4, 0, 4):
            for dz in (-4, 0, 4):
                cand.append(((x + dx + 32768) << 32) | ((y + dy + 32768) << 16) | (z + dz + 32768))
    assert torch.equal(prev.expand(3).keys, torch.unique(torch.cat(cand)))


def test_kernel_map_properties_full_size(pcc, level0):
    m = level0
    km = m.kernel_map(m, 3)
    nbr = km.natural_nbr()
    n = m.n
    rows = torch.arange(n, device=DEV, dtype=torch.int32)
    assert torch.equal(nbr[13], rows)                                   # centre tap
    for k in (0, 4, 9, 12):                                             # symmetry
        has = nbr[k] >= 0
        back = nbr[26 - k][nbr[k][has].long()]
        assert torch.equal(back, rows[has])
        assert int(has.sum()) == int((nbr[26 - k] >= 0).sum())
    # the tile-row permutation is a permutation and the tap mask covers every pair of its tile
    tiles = (n + 127) // 128
    pad = torch.full((27, tiles * 128 - n), -1, dtype=torch.int32, device=DEV)
    active = (torch.cat([km.nbr[:, :n], pad], 1).view(27, tiles, 128) >= 0).any(dim=2)
    assert km.n_groups == 1                                             # the input surface: mask-sorted tiles are >= 90 % useful (0.92)
    for g, (lo, hi) in enumerate(km.group_ranges()):
        assert torch.equal(torch.sort(km.order.view(km.n_groups, -1)[g].long()).values, torch.arange(n, device=DEV))
        tm = km.tapmask.view(km.n_groups, -1)[g].long()
        bits = ((tm.unsqueeze(0) >> torch.arange(27, device=DEV).unsqueeze(1)) & 1).bool()
        assert torch.equal(bits[lo:hi], active[lo:hi]) and not bool(bits[:lo].any()) and not bool(bits[hi:].any())


def test_automatic_offset_groups_full_size(pcc, level0):
    """pcc_kernel_map(n_groups = 0) on the surface levels of the frame: 0.92 useful -> one group; 0.86 and 0.76
    (249,574 and 68,423 rows) -> two; 17,957 rows cannot fill the GPU twice -> one (coords.cu, measured rule)."""
    m = level0
    want = [1, 2, 2, 1]
    for lv in range(4):
        km = m.kernel_map(m, 3)
        assert km.n_groups == want[lv], (lv, m.n, km.n_groups)
        assert km.order.numel() == km.n_groups * m.n and km.tapmask.numel() == km.n_groups * ((m.n + 127) // 128)
        m = m.stride2()


@pytest.mark.parametrize("cin,cout,mode", [(64, 64, 0), (128, 128, 0), (64, 4, 0), (128, 128, 1)])
def test_conv_scaling_property_full_size(pcc, level0, cin, cout, mode):
    m = level0
    g = torch.Generator(device=DEV).manual_seed(cin + cout)
    x = torch.randn((m.n, cin), device=DEV, generator=g)
    W = torch.randn((27, cin, cout), device=DEV, generator=g) / 20
    if mode == 1:  # tf32 operands: keep them representable so scaling by 2 stays exact
        x = (x.view(torch.int32) & ~0x1FFF).view(torch.float32)
        W = (W.view(torch.int32) & ~0x1FFF).view(torch.float32)
    km = m.kernel_map(m, 3, groups=1 if mode == 1 else None)
    a = pcc.ops.conv(x, km, W, None, mode=mode)
    b = pcc.ops.conv(2.0 * x, km, W, None, mode=mode)
    assert torch.isfinite(a).all()
    assert torch.equal((2.0 * a).view(torch.int32), b.view(torch.int32))
    # a row with a single neighbour (itself) under a centre-only kernel is a plain matrix product
    W1 = torch.zeros_like(W)
    W1[13] = W[13]
    c = pcc.ops.conv(x, km, W1, None, mode=mode)
    ref = x[:4096].double() @ W[13].double()
    assert torch.allclose(c[:4096].double(), ref, rtol=0, atol=2e-3 if mode else 1e-4)


def test_topk_property_full_size(pcc):
    g = torch.Generator(device=DEV).manual_seed(7)
    m = 3984493
    logits = torch.round(torch.randn((m, 4), device=DEV, generator=g) * 64) / 64   # many ties
    k = 799458
    sel = pcc.ops.topk_rows(logits, k).long()
    assert sel.shape[0] == k and bool((sel[1:] > sel[:-1]).all())
    chosen = torch.zeros(m, dtype=torch.bool, device=DEV)
    chosen[sel] = True
    v = logits[:, 0]
    thr = v[chosen].min()
    assert thr >= v[~chosen].max()
    # ties at the threshold go to the smaller rows
    tie_rows = torch.nonzero(v == thr).flatten()
    taken = chosen[tie_rows]
    n_taken = int(taken.sum())
    assert bool(taken[:n_taken].all()) and not bool(taken[n_taken:].any())


def _model(pcc, cfg):
    model = pcc.ColorModel(cfg)
    pcc.init_synthetic(model, 0)
    model.eval().to(DEV)
    model.update()
    return model


@pytest.mark.parametrize("config_name,q", [("OURS_CONFIG", (0.5, 0.5)), ("OURS_CONFIG", (0.0, 0.0)),
                                           ("OURS_CONFIG", (1.0, 1.0)), ("OURS_CONFIG", (0.0, 1.0)),
                                           ("OURS_CONFIG", (1.0, 0.0)), ("ABLATION_CONFIG", (0.5, 0.5)),
                                           # the rest of configs[1]'s sweep {0, 0.5, 1}^2
                                           ("OURS_CONFIG", (0.0, 0.5)), ("OURS_CONFIG", (0.5, 0.0)),
                                           ("OURS_CONFIG", (0.5, 1.0)), ("OURS_CONFIG", (1.0, 0.5))])
def test_codec_round_trip_full_size(pcc, frame, config_name, q):
    model = _model(pcc, getattr(pcc, config_name))
    x = frame
    N = x.shape[0]
    Q = pcc.qmap.q_map(x, *q)
    strings, shape, k, coords = model.compress(x.to(DEV), (Q[0].to(DEV), Q[1].to(DEV)))
    enc = model.entropy_model.last
    y_hat_enc = enc["y_symbols"].t().float() + enc["params"].F[:, enc["params"].F.shape[1] // 2:]
    assert k[-1] == [N] and len(k) == 3 and coords.shape[0] == model.last["y"].m.n
    out = model.decompress(coordinates=coords, strings=strings, shape=shape, k=k)
    # entropy round trip: the decoder's latent is the encoder's quantised latent, bit for bit
    assert torch.equal(model.last["y_hat"].F.view(torch.int32), y_hat_enc.view(torch.int32))
    # decoded cloud: N unique voxels, colours on the 8-bit grid in [0, 1]
    assert out.shape == (N, 6)
    keys, _ = pcc.ops.pack_coords(torch.cat([torch.zeros((N, 1), device=DEV), out[:, :3]], 1).int())
    assert torch.unique(keys).shape[0] == N
    rgb = out[:, 3:] * 255
    assert torch.allclose(rgb, torch.round(rgb), rtol=0, atol=1e-3) and float(rgb.min()) >= 0 and float(rgb.max()) <= 255.001
    # determinism, and independence of the input row order
    perm = torch.randperm(N, generator=torch.Generator().manual_seed(1))
    s2, sh2, k2, c2 = model.compress(x[perm].to(DEV), (Q[0][perm].to(DEV), Q[1][perm].to(DEV)))
    assert s2[0][0] == strings[0][0] and s2[1][0] == strings[1][0] and sh2 == shape and k2 == k
    assert torch.equal(c2, coords)


def test_quality_changes_the_stream_full_size(pcc, frame):
    """The quality map conditions the transforms: different q -> different latents and strings."""
    model = _model(pcc, pcc.OURS_CONFIG)
    xd = frame.to(DEV)
    sizes = {}
    for q in ((0.0, 0.0), (1.0, 1.0)):
        Q = pcc.qmap.q_map(frame, *q)
        s, sh, k, c = model.compress(xd, (Q[0].to(DEV), Q[1].to(DEV)))
        sizes[q] = (len(s[0][0]), s[0][0][:64])
    assert sizes[(0.0, 0.0)] != sizes[(1.0, 1.0)]
