"""Golden vectors produced by the reference's own model code (tests/golden/make_golden.py:
/root/reference/model/*.py executed unmodified over CPU stand-ins for MinkowskiEngine / CompressAI).
CPU: the oracle's restatement reproduces them bit for bit.  GPU: so does the product, through the C-ABI."""
import glob
import os

import numpy as np
import pytest
import torch

GOLDEN = sorted(p for p in glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz"))
                if not os.path.basename(p).startswith("train_"))  # train_*: tests/test_gpu_train.py


def _load(path):
    g = np.load(path)
    return {k: g[k] for k in g.files}


def _config(pcc, g):
    """The fixture's model config: a named product config, or (two-hyperprior case) the JSON the generator stored."""
    if "config_json" in g:
        import json
        return json.loads(str(g["config_json"]))
    return getattr(pcc, str(g["config"]))


def _strings_shape(g):
    """(strings, shape) as ColorModel.compress returns them (model/model.py:128-135)."""
    if "y_map_string" in g:
        return ([[[g["y_string"].tobytes()], [g["z_string"].tobytes()]],
                 [[g["y_map_string"].tobytes()], [g["z_map_string"].tobytes()]]],
                [[int(g["shape"][0])], [int(g["shape"][1])]])
    return [[g["y_string"].tobytes()], [g["z_string"].tobytes()]], g["shape"].tolist()


def _canon_out(out):
    key = (out[:, 0].astype(np.int64) << 40) | (out[:, 1].astype(np.int64) << 20) | out[:, 2].astype(np.int64)
    return out[np.argsort(key)]


def test_golden_fixtures_present():
    assert len(GOLDEN) >= 6 and any("twohyper" in p for p in GOLDEN)


def _qmap(pcc, g, x):
    """The fixture's quality map, rebuilt with the product's builders (pcc_b200.qmap) and checked
    against the features the generator stored."""
    kind = g["qkind"].tolist()
    qg, qa = float(g["q"][0]), float(g["q"][1])
    if not kind:
        qc, qf = pcc.qmap.q_map(x, qg, qa)
    elif kind[0] == 0.0:
        qc, qf = pcc.qmap.view_q_map(x, qg, qa, int(kind[1]), kind[2], kind[3])
    else:
        qc, qf = pcc.qmap.roi_q_map(x, qg, qa, int(kind[1]), kind[2])
    assert np.array_equal(qf.numpy().view(np.uint32), g["qf"].view(np.uint32))
    return qc, qf


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_reproduces_reference_code_outputs(pcc, oracle, path):
    g = _load(path)
    cfg = _config(pcc, g)
    model = pcc.ColorModel(cfg)
    pcc.init_synthetic(model, int(g["seed"]))
    oc = oracle.OracleCodec(pcc.export_params(model), cfg)
    x = torch.from_numpy(g["x"])
    qc, qf = _qmap(pcc, g, x)
    strings, shape, k, coords = oc.compress(g["x"], qc.numpy(), qf.numpy())
    want_strings, want_shape = _strings_shape(g)
    assert strings == want_strings, "bitstreams differ from the reference code's"
    assert shape == want_shape
    assert [int(kk[0]) for kk in k] == g["k"].tolist()
    assert np.array_equal(np.asarray(coords, np.int32), g["coordinates"])
    out = oc.decompress(coords, strings, shape, k)
    assert np.array_equal(_canon_out(np.asarray(out, np.float32)).view(np.uint32), g["out"].view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_product_reproduces_reference_code_outputs(pcc, path):
    g = _load(path)
    dev = "cuda:0"
    cfg = _config(pcc, g)
    model = pcc.ColorModel(cfg)
    pcc.init_synthetic(model, int(g["seed"]))
    model.eval().to(dev)
    model.update()
    x = torch.from_numpy(g["x"])
    qc, qf = _qmap(pcc, g, x)
    strings, shape, k, coords = model.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
    want_strings, want_shape = _strings_shape(g)
    assert strings == want_strings, "bitstreams differ from the reference code's"
    assert shape == want_shape
    assert [int(kk[0]) for kk in k] == g["k"].tolist()
    assert np.array_equal(coords.cpu().numpy(), g["coordinates"])
    # decode from the golden strings (not from what we just produced)
    out = model.decompress(coordinates=torch.from_numpy(g["coordinates"]), strings=want_strings, shape=want_shape,
                           k=[[int(v)] for v in g["k"]])
    assert np.array_equal(_canon_out(out.cpu().numpy()).view(np.uint32), g["out"].view(np.uint32))
