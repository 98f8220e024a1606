"""Checkpoints saved AFTER update() load into a fresh model (the reference saves its weights
after model.update(), train.py:291, and evaluate.py:81 loads that file; CompressAI handles the coder-table
buffers, registered empty, with update_registered_buffers)."""
import io

import numpy as np
import pytest
import torch


def _updated_state(pcc, cfg, seed):
    m = pcc.ColorModel(cfg)
    pcc.init_synthetic(m, seed)
    m.eval()
    m.update()
    buf = io.BytesIO()
    torch.save(m.state_dict(), buf)
    buf.seek(0)
    return m, torch.load(buf)


def test_post_update_checkpoint_loads_strict(pcc):
    src, sd = _updated_state(pcc, pcc.OURS_CONFIG, 3)
    assert sd["entropy_model.gaussian_conditional._quantized_cdf"].shape[0] == 64
    assert sd["entropy_model.entropy_bottleneck._offset"].shape == (128,)
    fresh = pcc.ColorModel(pcc.OURS_CONFIG)
    assert fresh.entropy_model.entropy_bottleneck._offset.numel() == 0
    missing, unexpected = fresh.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    for name in ("entropy_bottleneck", "gaussian_conditional"):
        a, b = getattr(src.entropy_model, name), getattr(fresh.entropy_model, name)
        for t in ("_quantized_cdf", "_cdf_length", "_offset"):
            assert torch.equal(getattr(a, t), getattr(b, t)), (name, t)
        for x, y in zip(a._tables(), b._tables()):   # host tables come from the LOADED buffers, no update() needed
            assert np.array_equal(x, y)
    assert torch.equal(src.entropy_model.gaussian_conditional.scale_table,
                       fresh.entropy_model.gaussian_conditional.scale_table)
    # a second load into an already-sized model, and a pre-update checkpoint, still work
    fresh.load_state_dict(sd, strict=True)
    pre = pcc.ColorModel(pcc.OURS_CONFIG)
    pcc.init_synthetic(pre, 3)
    fresh2 = pcc.ColorModel(pcc.OURS_CONFIG)
    fresh2.load_state_dict(pre.state_dict(), strict=True)
    assert fresh2.entropy_model.entropy_bottleneck._offset.numel() == 0


@pytest.mark.gpu
def test_post_update_checkpoint_codes_identically(pcc):
    dev = "cuda:0"
    src, sd = _updated_state(pcc, pcc.OURS_CONFIG, 5)
    x = pcc.synthetic.shell_frame(17.0, center=[64] * 3, grid=128, seed=5)
    qc, qf = pcc.synthetic.uniform_q_map(x, 0.3, 0.7)
    src.to(dev)
    want = src.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
    a = pcc.ColorModel(pcc.OURS_CONFIG)
    a.load_state_dict(sd)
    a.eval().to(dev)                       # loaded tables used as they are
    got = a.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
    assert got[0] == want[0] and got[1] == want[1] and got[2] == want[2] and torch.equal(got[3], want[3])
    a.update()                             # and after the reference's evaluate.py:83 model.update()
    again = a.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
    assert again[0] == want[0]
    out = a.decompress(coordinates=again[3], strings=again[0], shape=again[1], k=again[2])
    ref = src.decompress(coordinates=want[3], strings=want[0], shape=want[1], k=want[2])
    assert torch.equal(out, ref)
