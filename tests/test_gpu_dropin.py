"""INTEGRATION.md "Option A", demonstrated: the REFERENCE'S OWN model code — model/{model,transforms,blocks,
entropy_models}.py and utils.py, compiled unmodified to byte code by oracle/build_ref.py (the reference tree
itself does not exist on the GPU box) — runs on the GPU over THIS repo's module surface:

    import MinkowskiEngine as ME               ->  pcc_b200.me
    compressai.entropy_models / .models.base   ->  pcc_b200.compressai_compat
    compressai.layers.GDN                      ->  pcc_b200.compressai_compat.GDN

and must reproduce, bit for bit, the five golden fixtures the same reference code produced over the CPU
stand-ins (tests/golden/make_golden.py).  This is the one test that drives `SparseTensor(coordinates=float…)`,
`coordinate_manager`, `features_at_coordinates`, `MinkowskiPruning`, the `Sorted*` subclasses
(model/entropy_models.py:12-102, `super().forward`) and `EntropyBottleneck/GaussianConditional
.compress/.decompress` through the reference's own call sequence (utils.py:436-461, model/model.py:95-208).
"""
import glob
import importlib
import os
import sys
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = sorted(p for p in glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz"))
                if not os.path.basename(p).startswith("train_"))
_NAMES = ("MinkowskiEngine", "compressai", "compressai.entropy_models", "compressai.models",
          "compressai.models.base", "compressai.layers", "open3d", "bitstream")


@pytest.fixture()
def reference_over_product(pcc):
    """sys.modules as a maintainer following INTEGRATION.md §2 would leave it, undone afterwards."""
    from oracle import build_ref
    if not build_ref.available():
        build_ref.build()                       # builder container: /root/reference is there
    if not build_ref.available():
        have = sorted(os.listdir(build_ref.OUT)) if os.path.isdir(build_ref.OUT) else None
        pytest.fail("oracle/_ref/pyref is incomplete: run `python __graft_entry__.py build` where /root/reference "
                    f"exists (found {have})")
    saved = {n: sys.modules.get(n) for n in _NAMES}
    finder = build_ref.ReferenceFinder()
    cc = pcc.compressai_compat
    ca = types.ModuleType("compressai")
    ca_em = types.ModuleType("compressai.entropy_models")
    ca_em.EntropyBottleneck, ca_em.GaussianConditional = cc.EntropyBottleneck, cc.GaussianConditional
    ca_models = types.ModuleType("compressai.models")
    ca_base = types.ModuleType("compressai.models.base")
    ca_base.CompressionModel = cc.CompressionModel
    ca_layers = types.ModuleType("compressai.layers")
    ca_layers.GDN = cc.GDN
    ca.entropy_models, ca.models, ca.layers, ca_models.base = ca_em, ca_models, ca_layers, ca_base
    o3d = types.ModuleType("open3d")            # imported at model/model.py:8 and utils.py:2, unused on the path
    bs = types.ModuleType("bitstream")          # model/model.py:9, only the file container uses it
    bs.BitStream = type("BitStream", (), {})
    for name, mod in zip(_NAMES, (pcc.me, ca, ca_em, ca_models, ca_base, ca_layers, o3d, bs)):
        sys.modules[name] = mod
    sys.meta_path.insert(0, finder)
    for m in [m for m in sys.modules if m == "model" or m.startswith("model.") or m == "utils"]:
        del sys.modules[m]
    try:
        mod = importlib.import_module("model.model")
        assert mod.__spec__.origin.endswith("model.model.bc") and build_ref.OUT in mod.__spec__.origin
        assert mod.ME is pcc.me
        yield mod
    finally:
        sys.meta_path.remove(finder)
        for m in [m for m in sys.modules if m == "model" or m.startswith("model.") or m == "utils"]:
            del sys.modules[m]
        for n, v in saved.items():
            if v is None:
                sys.modules.pop(n, None)
            else:
                sys.modules[n] = v


def _canon_out(out):
    key = (out[:, 0].astype(np.int64) << 40) | (out[:, 1].astype(np.int64) << 20) | out[:, 2].astype(np.int64)
    return out[np.argsort(key)]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_reference_model_code_runs_over_product_modules(pcc, reference_over_product, path):
    from test_golden import _config, _strings_shape
    g = np.load(path)
    cfg = _config(pcc, g)
    want_strings, want_shape = _strings_shape(g)
    ref_model = reference_over_product.ColorModel(cfg)           # the reference's class, the product's layers
    assert type(ref_model).__module__ == "model.model"
    assert isinstance(ref_model.g_a.down_1, pcc.me.MinkowskiConvolution)
    ours = pcc.ColorModel(cfg)
    pcc.init_synthetic(ours, int(g["seed"]))
    missing, unexpected = ref_model.load_state_dict(ours.state_dict(), strict=True)   # identical key sets
    assert not missing and not unexpected
    ref_model.eval().to(DEV)
    ref_model.update()                                             # evaluate.py:83
    x = torch.from_numpy(g["x"]).to(DEV)
    n = x.shape[0]
    ME = pcc.me
    with torch.no_grad():
        # utils.py:436-445 (non-uniform branch for the view / ROI fixtures)
        Q = ME.SparseTensor(coordinates=torch.cat([torch.zeros((n, 1), device=DEV), x[:, :3]], dim=1),
                            features=torch.from_numpy(g["qf"]).to(DEV), device=x.device)
        strings, shape, k, coordinates = ref_model.compress(x, Q)                      # utils.py:452
        assert strings == want_strings, "bitstreams differ from the fixture"
        assert shape == want_shape
        assert [int(kk[0]) for kk in k] == g["k"].tolist()
        c = coordinates.cpu().numpy().astype(np.int32)
        ck = (c[:, 1].astype(np.int64) << 40) | (c[:, 2].astype(np.int64) << 20) | c[:, 3].astype(np.int64)
        assert np.array_equal(c[np.argsort(ck)], g["coordinates"])
        out = ref_model.decompress(coordinates=coordinates.clone(), strings=strings, shape=shape, k=k)  # utils.py:461
    got, want = _canon_out(out.cpu().numpy().astype(np.float32)), g["out"]
    assert np.array_equal(got[:, :3], want[:, :3]), "decoded occupancy differs from the fixture"
    # The last statement of ColorModel.decompress (model/model.py:206) is the reference's own torch expression
    # `clamp(round(F * 255), 0, 255) / 255`, here evaluated on a CUDA tensor: torch's CUDA division by a Python
    # scalar multiplies by the reciprocal (ATen BinaryDivTrueKernel.cu), the CPU run that made the fixture
    # divided.  So the 8-bit colour codes must be identical and the floats within one unit in the last place.
    assert np.array_equal(np.rint(got[:, 3:] * 255.0), np.rint(want[:, 3:] * 255.0)), "decoded colours differ"
    gc_, wc_ = got[:, 3:] + np.float32(0), want[:, 3:] + np.float32(0)   # -0.0 (clamp of a rounded small negative) == +0.0
    ulp = np.abs(gc_.view(np.int32).astype(np.int64) - wc_.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1
