#!/usr/bin/env python
"""Generates tests/golden/train_*.npz by running the REFERENCE'S OWN training forward and losses.

  python tests/golden/make_golden_train.py     (in the builder container: needs /root/reference)

/root/reference/model/*.py (``ColorModel.forward``, model/model.py:51-93) and /root/reference/loss.py
(``Loss``: Multiscale_FocalLoss + ColorLoss + BPPLoss x2, the ``loss:`` section of configs/Ours.yaml)
are imported unmodified and run in train() mode over the differentiable CPU stand-ins of
oracle/train_shims.py; ``loss.backward()`` and ``aux_loss().backward()`` (train.py:189-203) then give
the gradient of every parameter.  The fixture holds the batch, the noise draws, the loss terms and a
fingerprint of every gradient (norm + 16 sampled entries); the product test replays the noise and
must reproduce them within the stated fp32 tolerance (tests/test_gpu_train.py).

The tape runs in float64 (``torch.set_default_dtype``) on the fp32-valued weights, inputs and noise:
the fixture is the gradient of the function itself, not of one fp32 evaluation order.  A second pass
with a float32 tape is stored beside it as ``gdev32/<name>``: how far ONE float32 evaluation of the same
code lands from the float64 gradient (a few bias sums of the 2-channel quality branch cancel to 0.1-0.6 %
of their terms) — the yardstick for the product's tolerance.
"""
import os
import sys

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

CASES = [  # name, yaml, product config, [(radius, centre, colour seed)] per batch element, weight seed
    ("train_ours_b2", "configs/Ours.yaml", "OURS_CONFIG", [(9.0, 32, 0), (7.0, 30, 1)], 0),
    ("train_ablation_b1", "configs/Ablation_NoCondition_Convolution.yaml", "ABLATION_CONFIG", [(10.0, 32, 2)], 1),
]


def make_batch(pcc_b200, clouds):
    """-> coords [N,4] int32 (batch first), colours [N,3], q [N,2] in [0,1] (per-cloud gradient + constant)."""
    cs, fs, qs = [], [], []
    for b, (radius, centre, seed) in enumerate(clouds):
        x = pcc_b200.synthetic.shell_frame(radius, center=[centre] * 3, grid=64, seed=seed)
        n = x.shape[0]
        cs.append(torch.cat([torch.full((n, 1), b, dtype=torch.int32), x[:, :3].int()], dim=1))
        fs.append(x[:, 3:6].float())
        ramp = ((x[:, 1] - x[:, 1].min()) / max(float(x[:, 1].max() - x[:, 1].min()), 1.0)).float()
        qs.append(torch.stack([ramp, torch.full((n,), 0.25 + 0.5 * b)], dim=1))
    return torch.cat(cs), torch.cat(fs), torch.cat(qs)


def fingerprint(g, rng):
    g = g.detach().double().reshape(-1)
    idx = rng.integers(0, g.numel(), size=16)
    return float(g.norm()), idx.astype(np.int64), g[idx].numpy()


def run_case(pcc_b200, TS, case, loss_cfg, q_cfg, dtype):
    """One pass of the reference's forward + losses + backward with the tape in `dtype`."""
    name, yml, cfg_name, clouds, wseed = case
    torch.set_default_dtype(dtype)
    cfg = yaml.safe_load(open(os.path.join(REF, yml)))["model"]
    assert cfg == getattr(pcc_b200, cfg_name)
    ref_model, ref_loss = TS.load_reference(cfg, REF)
    ours = pcc_b200.ColorModel(cfg)
    pcc_b200.init_synthetic(ours, wseed)
    missing, unexpected = ref_model.load_state_dict(dict(ours.state_dict()), strict=False)
    assert not missing, missing
    ref_model.train()
    coords, colours, q = make_batch(pcc_b200, clouds)
    colours, q = colours.float(), q.float()
    lam = pcc_b200.qmap.lambda_map(q, q_cfg)           # data/q_map.py:188-205 (exponential mode)
    import MinkowskiEngine as ME  # the differentiable stand-in
    x = ME.SparseTensor(coordinates=coords, features=colours)
    Q = ME.SparseTensor(coordinates=coords, features=q)
    Lambda = ME.SparseTensor(coordinates=coords, features=lam)
    TS.reseed(1234 + wseed)
    out = ref_model(x, Q, Lambda)                      # model/model.py:51-93
    loss_fn = ref_loss.Loss({k: dict(v) for k, v in loss_cfg.items()})
    total, parts = loss_fn(x, out)                     # loss.py:41-66
    total.backward()                                   # train.py:191
    aux = ref_model.aux_loss()
    aux.backward()                                     # train.py:201
    rng = np.random.default_rng(7)
    rec = {"config": cfg_name, "wseed": wseed, "coords": coords.numpy(), "colours": colours.numpy(),
           "q": q.numpy(), "lam": lam.numpy(), "total": float(total.detach()), "aux": float(aux.detach()),
           "n_noise": len(TS.noise_log)}
    for i, nz in enumerate(TS.noise_log):
        rec[f"noise{i}"] = nz.numpy()
    for k, v in parts.items():
        rec["loss/" + k] = float(v.detach())
    names, seen = [], set()
    for pname, p in ref_model.named_parameters():
        if id(p) in seen or p.grad is None:
            continue
        seen.add(id(p))
        norm, idx, vals = fingerprint(p.grad, rng)
        names.append(pname)
        rec["gnorm/" + pname], rec["gidx/" + pname], rec["gval/" + pname] = norm, idx, vals
    rec["grad_names"] = np.array(names)
    return rec


def main():
    import pcc_b200
    from oracle import train_shims as TS
    full = yaml.safe_load(open(os.path.join(REF, "configs/Ours.yaml")))
    loss_cfg, q_cfg = full["loss"], full["q_map"]
    for case in CASES:
        name = case[0]
        rec = run_case(pcc_b200, TS, case, loss_cfg, q_cfg, torch.float64)
        r32 = run_case(pcc_b200, TS, case, loss_cfg, q_cfg, torch.float32)
        torch.set_default_dtype(torch.float32)
        names = rec["grad_names"].tolist()
        assert names == r32["grad_names"].tolist()
        worst = []
        for n in names:  # how far ONE float32 evaluation of the same code is from the float64 gradient
            ref = max(rec["gnorm/" + n], 1e-30)
            dev = max(np.abs(r32["gval/" + n] - rec["gval/" + n]).max(), abs(r32["gnorm/" + n] - rec["gnorm/" + n])) / ref
            rec["gdev32/" + n] = dev
            worst.append((dev, n))
        worst.sort(reverse=True)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **rec)
        print(f"{name}: {rec['coords'].shape[0]} points, total {rec['total']:.6f} (float32 tape {r32['total']:.6f}) "
              f"aux {rec['aux']:.4f}, parts { {k[5:]: round(v, 6) for k, v in rec.items() if k.startswith('loss/')} }, "
              f"{len(names)} gradients, {rec['n_noise']} noise draws; float32-tape deviation of the worst gradients: "
              + ", ".join(f"{n} {e:.1e}" for e, n in worst[:4]))


if __name__ == "__main__":
    main()
