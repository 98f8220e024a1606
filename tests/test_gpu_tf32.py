"""Tensor-core mode (PCC_MODE_TF32, tcgen05.mma kind::tf32) against the exact fp32 mode.

Floating-point tolerance (stated here, as north_star asks): operands are read as tf32 (10-bit
mantissa, truncation), accumulation is fp32, so one convolution differs from the exact mode by at
most 2e-3 * max|out| (observed ~3e-4); the end-to-end bound on latents / attributes is asserted in
test_codec_tf32_*.  Integer outputs (coordinates, kernel maps) are mode-independent."""
import numpy as np
import pytest
import torch

from conftest import random_cloud

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("cin,cout", [(32, 32), (64, 64), (128, 128), (128, 64), (64, 128), (128, 256),
                                      (192, 256), (128, 192), (64, 32), (32, 96)])
@pytest.mark.parametrize("n", [100, 5000])
def test_conv_tf32_close_to_exact(pcc, cin, cout, n):
    ops = pcc.ops
    rng = np.random.default_rng(cin + cout + n)
    c = random_cloud(n, 12 if n < 1000 else 26, 3, 1, -2)
    m, _ = ops.canonical_map(torch.from_numpy(c).to(DEV))
    x = torch.from_numpy(rng.standard_normal((m.n, cin)).astype(np.float32)).to(DEV)
    W = torch.from_numpy((rng.standard_normal((27, cin, cout)) / np.sqrt(cin * 9)).astype(np.float32)).to(DEV)
    b = torch.from_numpy(rng.standard_normal((1, cout)).astype(np.float32)).to(DEV)
    film = torch.from_numpy(rng.standard_normal((m.n, 2 * cout)).astype(np.float32)).to(DEV)
    res = torch.from_numpy(rng.standard_normal((m.n, cout)).astype(np.float32)).to(DEV)
    km = m.kernel_map(m, 3, groups=1)
    for kw in ({}, {"act": ops.ACT_RELU, "film": film, "residual": res}, {"act": ops.ACT_LEAKY}):
        ref = ops.conv(x, km, W, b, mode=ops.MODE_EXACT, **kw)
        got = ops.conv(x, km, W, b, mode=ops.MODE_TF32, **kw)
        err = (got - ref).abs().max().item()
        scale = ref.abs().max().item()
        assert err <= 2e-3 * scale, (err, scale, kw.keys())
    # exactly representable operands -> tf32 products and fp32 sums are exact for small integers
    xi = torch.randint(-4, 5, (m.n, cin), device=DEV).float()
    Wi = torch.randint(-2, 3, (27, cin, cout), device=DEV).float()
    assert torch.equal(ops.conv(xi, km, Wi, None, mode=ops.MODE_TF32), ops.conv(xi, km, Wi, None, mode=ops.MODE_EXACT))


@pytest.mark.parametrize("ks,transpose", [(3, True), (2, True), (3, False)])
def test_conv_tf32_strided_and_transposed(pcc, ks, transpose):
    ops = pcc.ops
    c = random_cloud(4000, 24, 5, 1, 0)
    c[:, 1:] *= 4
    m, _ = ops.canonical_map(torch.from_numpy(c).to(DEV), 4)
    out_map = m.expand(ks) if transpose else m.stride2()
    km = m.kernel_map(out_map, ks, transpose=transpose, groups=1)
    xi = torch.randint(-4, 5, (m.n, 64), device=DEV).float()
    Wi = torch.randint(-2, 3, (ks ** 3, 64, 128), device=DEV).float()
    assert torch.equal(ops.conv(xi, km, Wi, None, mode=ops.MODE_TF32), ops.conv(xi, km, Wi, None, mode=ops.MODE_EXACT))


def test_tf32_small_channels_take_exact_path(pcc):
    """Stated policy: layers the tensor-core path cannot tile (cin or cout not a multiple of 32)
    run the exact CUDA-core kernel even in tf32 mode -> bit-identical."""
    ops = pcc.ops
    c = random_cloud(800, 14, 7, 1, 0)
    m, _ = ops.canonical_map(torch.from_numpy(c).to(DEV))
    x = torch.randn((m.n, 4), device=DEV)
    W = torch.randn((27, 4, 64), device=DEV)
    km = m.kernel_map(m, 3, groups=1)
    ref = ops.conv(x, km, W, None, mode=ops.MODE_EXACT)
    assert torch.equal(ops.conv(x, km, W, None, mode=ops.MODE_TF32, round_out=False), ref)
    # by default tf32 mode stores activations rounded to tf32 (round-to-nearest, unbiased), so that
    # the tensor core's operand truncation downstream is a no-op
    got = ops.conv(x, km, W, None, mode=ops.MODE_TF32)
    assert (got - ref).abs().max().item() <= 2.0 ** -11 * ref.abs().max().item() * 1.001
    assert torch.equal(got.view(torch.int32) & 0x1FFF, torch.zeros_like(got, dtype=torch.int32))


def _codec(pcc, mode, radius=26.0, seed=0):
    ops = pcc.ops
    ops.set_default_mode(mode)
    try:
        model = pcc.ColorModel(pcc.OURS_CONFIG)
        pcc.codec.init_synthetic(model, seed)
        model.eval().to(DEV)
        model.update()
        x = pcc.synthetic.shell_frame(radius, center=[64, 64, 64], grid=128, seed=seed)
        qc, qf = pcc.synthetic.uniform_q_map(x)
        strings, shape, k, coords = model.compress(x.to(DEV), (qc.to(DEV), qf.to(DEV)))
        y = model.last["y"].F.clone()
        enc = {kk: (v.F.clone() if hasattr(v, "F") else v.clone()) for kk, v in model.entropy_model.last.items()}
        out = model.decompress(coordinates=coords, strings=strings, shape=shape, k=k)
        return {"strings": strings, "shape": shape, "k": k, "coords": coords, "y": y, "enc": enc, "out": out,
                "y_hat": model.last["y_hat"].F.clone()}
    finally:
        ops.set_default_mode(ops.MODE_EXACT)


def test_codec_tf32_roundtrip_and_tolerance(pcc):
    ops = pcc.ops
    ex = _codec(pcc, ops.MODE_EXACT)
    tc = _codec(pcc, ops.MODE_TF32)
    # integer structure is mode-independent
    assert tc["shape"] == ex["shape"] and tc["k"] == ex["k"]
    assert torch.equal(tc["coords"], ex["coords"])
    # latent features within the stated tolerance: max-abs 1e-3 relative to the latent's scale
    scale = ex["y"].abs().max().item()
    err = (tc["y"] - ex["y"]).abs().max().item()
    print(f"tf32 latent max-abs err {err:.3e} on scale {scale:.3f} (rel {err / scale:.2e})")
    assert err <= 4e-3 * max(scale, 1.0)
    # decode round trip in tf32 mode: the decoder (same kernels) recovers exactly the encoder's
    # quantised latent, so the bitstream is decodable without drift
    y_hat_enc = tc["enc"]["y_symbols"].t().float() + tc["enc"]["params"][:, 128:]
    assert torch.equal(tc["y_hat"], y_hat_enc)
    assert tc["out"].shape == ex["out"].shape
    # discrete agreement with the exact mode (reported, loosely bounded: tf32 may flip symbols at
    # rounding boundaries and voxels at the top-k threshold)
    sym_mismatch = (tc["enc"]["y_symbols"] != ex["enc"]["y_symbols"]).float().mean().item()
    print(f"tf32 symbol mismatch rate vs exact {sym_mismatch:.4f}")
    assert sym_mismatch < 0.02, sym_mismatch
    a = {tuple(r) for r in tc["out"][:, :3].int().cpu().tolist()}
    b = {tuple(r) for r in ex["out"][:, :3].int().cpu().tolist()}
    iou = len(a & b) / len(a | b)
    print(f"tf32 occupancy IoU vs exact {iou:.4f}")
    # reported, not a parity claim: with random-init weights the occupancy logits are nearly flat, so
    # a 1e-3 perturbation reorders many voxels at the three cascaded top-k thresholds (measured 0.54
    # on this frame).  The bit-exact occupancy claim belongs to the exact mode (test_gpu_codec.py).
    assert iou > 0.25, iou


@pytest.mark.parametrize("cin", [64, 128, 20])
def test_factorised_head_matches_direct_convolution(pcc, cin):
    """pcc_conv_head0 (projection, then a scalar gather) == channel 0 of the exact convolution up to the
    fp32 summation order: max error 1e-5 of the largest logit."""
    import numpy as np
    from conftest import random_cloud
    rng = np.random.default_rng(cin)
    c = random_cloud(5000, 24, cin, 2, 0)
    m, _ = pcc.ops.canonical_map(torch.from_numpy(c).to("cuda:0"), 1)
    x = torch.from_numpy(rng.standard_normal((m.n, cin)).astype(np.float32)).to("cuda:0")
    W = torch.from_numpy((rng.standard_normal((27, cin, 7)) / np.sqrt(cin * 9)).astype(np.float32)).to("cuda:0")
    b = torch.from_numpy(rng.standard_normal((1, 7)).astype(np.float32)).to("cuda:0")
    pcc.ops.set_tap_groups(1)
    try:
        km = m.kernel_map(m, 3)
        ref = pcc.ops.conv(x, km, W, b, mode=pcc.ops.MODE_EXACT)[:, 0]
        got = pcc.ops.conv_head0(x, km, W, b)[:, 0]
    finally:
        pcc.ops.set_tap_groups(None)
    err = (got - ref).abs().max().item() / ref.abs().max().item()
    assert err <= 1e-5, err
