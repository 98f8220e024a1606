"""CPU pins of the training oracle (oracle/train_shims.py) and of the host-side training logic.

The differentiable stand-ins are checked against independent torch operators: the sparse
convolution and its gradients equal ``conv3d`` / ``conv_transpose3d`` autograd on the zero-filled
dense grid; LowerBound follows CompressAI's published gradient rule; the training-mode Gaussian
likelihood integrates to one over the integer grid.  When /root/reference is present (builder
container) the committed fixture is regenerated from the reference's own code and compared.
"""
import os

import numpy as np
import pytest
import torch

from conftest import random_cloud

HERE = os.path.dirname(os.path.abspath(__file__))


def _put(d, c, feats):
    d = d.clone()
    d[0, :, c[:, 0], c[:, 1], c[:, 2]] = feats.t()
    return d


def test_sparse_conv_and_gradients_equal_dense_conv3d(oracle):
    from oracle import train_shims as TS
    P = oracle.prims
    rng = np.random.default_rng(0)
    c = random_cloud(300, 10, 1, 1, 0)
    keys = np.sort(P.pack_keys(c))
    cs = P.unpack_keys(keys)
    cin, cout = 3, 5
    x = torch.from_numpy(rng.standard_normal((len(cs), cin))).double().requires_grad_(True)
    W = torch.from_numpy(rng.standard_normal((27, cin, cout))).double().requires_grad_(True)
    nbr = P.kernel_map(keys, cs, P.kernel_offsets(3))
    out = TS._conv_apply(x, nbr, W, None, len(cs))
    G = torch.from_numpy(rng.standard_normal(out.shape)).double()
    out.backward(G)
    # dense: ME kernel index k = dx + 3 dy + 9 dz (first axis fastest) -> conv3d weight [co, ci, x, y, z]
    xd = x.detach().clone().requires_grad_(True)
    Wd = W.detach().clone().requires_grad_(True)
    dense = _put(torch.zeros((1, cin, 12, 12, 12), dtype=torch.float64), torch.from_numpy(cs[:, 1:].astype(np.int64) + 1), xd)
    w5 = Wd.reshape(3, 3, 3, cin, cout).permute(4, 3, 2, 1, 0)  # [dz,dy,dx,ci,co] -> [co,ci,dx,dy,dz]
    full = torch.nn.functional.conv3d(dense, w5, padding=1)
    ci = torch.from_numpy(cs[:, 1:].astype(np.int64) + 1)
    ref = full[0, :, ci[:, 0], ci[:, 1], ci[:, 2]].t()
    ref.backward(G)
    assert torch.allclose(out, ref, atol=1e-10)
    assert torch.allclose(x.grad, xd.grad, atol=1e-10)
    assert torch.allclose(W.grad, Wd.grad, atol=1e-10)


def test_generative_transpose_and_gradients_equal_dense(oracle):
    from oracle import train_shims as TS
    P = oracle.prims
    rng = np.random.default_rng(1)
    c = random_cloud(60, 5, 2, 1, 0)
    c[:, 1:] *= 2
    keys = np.sort(P.pack_keys(c))
    cs = P.unpack_keys(keys)
    cin, cout = 2, 3
    x = torch.from_numpy(rng.standard_normal((len(cs), cin))).double().requires_grad_(True)
    W = torch.from_numpy(rng.standard_normal((8, cin, cout))).double().requires_grad_(True)
    out_coords = P.expand_coords(cs, 2, 1)
    nbr = P.kernel_map(keys, out_coords, -P.kernel_offsets(2))
    out = TS._conv_apply(x, nbr, W, None, len(out_coords))
    G = torch.from_numpy(rng.standard_normal(out.shape)).double()
    out.backward(G)
    xd, Wd = x.detach().clone().requires_grad_(True), W.detach().clone().requires_grad_(True)
    dense = _put(torch.zeros((1, cin, 5, 5, 5), dtype=torch.float64), torch.from_numpy(cs[:, 1:].astype(np.int64) // 2), xd)
    w5 = Wd.reshape(2, 2, 2, cin, cout).permute(3, 4, 2, 1, 0)  # conv_transpose3d weight [ci, co, x, y, z]
    full = torch.nn.functional.conv_transpose3d(dense, w5, stride=2)
    oc = torch.from_numpy(out_coords[:, 1:].astype(np.int64))
    ref = full[0, :, oc[:, 0], oc[:, 1], oc[:, 2]].t()
    ref.backward(G)
    assert torch.allclose(out, ref, atol=1e-10)
    assert torch.allclose(x.grad, xd.grad, atol=1e-10) and torch.allclose(W.grad, Wd.grad, atol=1e-10)


def test_lower_bound_gradient_rule():
    from oracle import train_shims as TS
    x = torch.tensor([0.05, 0.05, 0.2, 0.2], requires_grad=True)
    y = TS._LowerBoundFn.apply(x, torch.tensor([0.11]))
    y.backward(torch.tensor([1.0, -1.0, 1.0, -1.0]))
    assert y.tolist() == pytest.approx([0.11, 0.11, 0.2, 0.2])
    assert x.grad.tolist() == [0.0, -1.0, 1.0, -1.0]  # below the bound only a gradient that raises x passes


def test_training_likelihood_sums_to_one_over_the_integer_grid():
    from oracle import train_shims as TS
    gc = TS.GaussianConditional(None).train()
    v = torch.arange(-40, 41, dtype=torch.float32).reshape(1, 1, -1) + 0.3
    TS.reseed(0)
    _, lik = gc(v, torch.full_like(v, 2.5), means=torch.full_like(v, 0.1))
    noise = TS.noise_log[0]
    # the noisy points are no longer on a unit grid; evaluate the density on the exact shifted grid instead
    s, m = 2.5, 0.1
    val = (v - m).abs()
    cum = lambda t: 0.5 * torch.erfc(-(2 ** -0.5) * t)
    assert float((cum((0.5 - val) / s) - cum((-0.5 - val) / s)).sum()) == pytest.approx(1.0, abs=1e-5)
    assert noise.abs().max() <= 0.5 and lik.min() >= 1e-9


def test_lambda_map_matches_reference_formula(pcc):
    q = torch.tensor([[0.0, 0.0], [1.0, 1.0], [0.5, 0.25]])
    lam = pcc.qmap.lambda_map(q)
    assert lam[0].tolist() == [0.0, 0.0]                       # 2^0 - 1
    assert lam[1].tolist() == pytest.approx([199.0, 6399.0], rel=1e-5)
    assert float(lam[2, 0]) == pytest.approx(200 ** 0.5 - 1, rel=1e-5)
    assert float(lam[2, 1]) == pytest.approx(6400 ** 0.25 - 1, rel=1e-5)


def test_product_training_modules_refuse_cpu(pcc):
    """No CPU fallback on the training path either."""
    with pytest.raises(pcc.PccError):
        pcc.ops.conv_wgrad(torch.zeros(4, 4), torch.zeros(4, 4), None, 4, 4)


@pytest.mark.skipif(not os.path.isdir("/root/reference/model"), reason="reference sources not present")
def test_fixture_regenerates_from_reference_code(pcc, oracle):
    """The committed fixture is what the reference's own forward + loss give today (ablation case, 1250 points)."""
    import sys
    import yaml
    from oracle import train_shims as TS
    sys.path.insert(0, os.path.join(HERE, "golden"))
    g = np.load(os.path.join(HERE, "golden", "train_ablation_b1.npz"))
    full = yaml.safe_load(open("/root/reference/configs/Ours.yaml"))
    cfg = yaml.safe_load(open("/root/reference/configs/Ablation_NoCondition_Convolution.yaml"))["model"]
    saved = {k: sys.modules.get(k) for k in TS.NAMES}
    torch.set_default_dtype(torch.float64)  # the fixture's tape is float64
    try:
        ref_model, ref_loss = TS.load_reference(cfg)
        ours = pcc.ColorModel(cfg)
        pcc.init_synthetic(ours, int(g["wseed"]))
        ref_model.load_state_dict(dict(ours.state_dict()), strict=False)
        ref_model.train()
        import MinkowskiEngine as ME
        coords = torch.from_numpy(g["coords"])
        x = ME.SparseTensor(coordinates=coords, features=torch.from_numpy(g["colours"]))
        Q = ME.SparseTensor(coordinates=coords, features=torch.from_numpy(g["q"]))
        Lam = ME.SparseTensor(coordinates=coords, features=torch.from_numpy(g["lam"]))
        TS.reseed(1234 + int(g["wseed"]))
        total, parts = ref_loss.Loss({k: dict(v) for k, v in full["loss"].items()})(x, ref_model(x, Q, Lam))
        assert float(total.detach()) == pytest.approx(float(g["total"]), rel=1e-9)
        for k, v in parts.items():
            assert float(v.detach()) == pytest.approx(float(g["loss/" + k]), rel=1e-9)
    finally:
        torch.set_default_dtype(torch.float32)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        for m in [m for m in sys.modules if m == "model" or m.startswith("model.") or m in ("utils", "loss")]:
            del sys.modules[m]
