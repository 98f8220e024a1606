#!/usr/bin/env python
"""Useful-MAC fraction of every conv launch of one encode+decode at tile (128-row) and at
32-row-group granularity: pairs / (rows_per_unit * sum over units of active taps).  GPU tool."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pcc_b200
from pcc_b200 import ops
dev = "cuda:0"
model = pcc_b200.ColorModel(pcc_b200.OURS_CONFIG); pcc_b200.init_synthetic(model, 0); model.eval().to(dev); model.update()
x = pcc_b200.synthetic.shell_frame(252.0, seed=0); qc, qf = pcc_b200.synthetic.uniform_q_map(x, 0.5, 0.5)
seen, rows = {}, []
orig = ops.conv
def hook(feats, kmap, weight, *a, **k):
    w = weight if weight.dim() == 3 else weight.unsqueeze(0)
    rows.append((kmap, w.shape[1], w.shape[2]))
    return orig(feats, kmap, weight, *a, **k)
ops.conv = hook
pcc_b200.codec.ops.conv = hook
s, sh, k, c = model.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
model.decompress(coordinates=c, strings=s, shape=sh, k=k)
def eff(kmap, g):
    n = kmap.n_out
    if kmap.K == 1 or n == 0: return 1.0, n
    nb = kmap.nbr[:, :n] >= 0
    pairs = int(nb.sum())
    pad = (-n) % g
    if pad: nb = torch.cat([nb, torch.zeros((kmap.K, pad), dtype=torch.bool, device=nb.device)], 1)
    act = nb.view(kmap.K, -1, g).any(dim=2).sum()
    return pairs / (g * int(act)), pairs
tot = {128: 0.0, 32: 0.0, "pairs": 0.0}
print("cin cout K n_out eff128 eff32")
for kmap, cin, cout in rows:
    if id(kmap) not in seen:
        seen[id(kmap)] = (eff(kmap, 128), eff(kmap, 32))
    (e128, pairs), (e32, _) = seen[id(kmap)]
    if cin >= 64 and cout >= 64:
        fl = pairs * cin * cout
        tot["pairs"] += fl; tot[128] += fl / e128; tot[32] += fl / e32
    print(cin, cout, kmap.K, kmap.n_out, round(e128, 3), round(e32, 3))
print("big layers: useful/issued at 128-row tiles %.3f, at 32-row groups %.3f -> issued work ratio %.3f" % (
    tot["pairs"] / tot[128], tot["pairs"] / tot[32], tot[32] / tot[128]))
