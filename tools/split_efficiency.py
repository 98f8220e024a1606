#!/usr/bin/env python
"""Useful-MAC fraction of the conv launches of one encode+decode when the 27 kernel offsets are cut
into S groups, each with its own mask-sorted tile rows (S passes chained through the accumulator)."""
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
s, sh, k, c = model.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
model.decompress(coordinates=c, strings=s, shape=sh, k=k)
def eff(kmap, S):
    n = kmap.n_out
    nb = kmap.natural_nbr() >= 0
    pairs = int(nb.sum())
    per = (kmap.K + S - 1) // S
    issued = 0
    for g in range(S):
        sub = nb[g * per:(g + 1) * per]
        if sub.shape[0] == 0: continue
        w = (1 << torch.arange(sub.shape[0], device=dev)).unsqueeze(1)
        m = (sub.long() * w).sum(0)
        o = torch.argsort(m)
        sub = sub[:, o]
        pad = (-n) % 128
        if pad: sub = torch.cat([sub, torch.zeros((sub.shape[0], pad), dtype=torch.bool, device=dev)], 1)
        issued += 128 * int(sub.view(sub.shape[0], -1, 128).any(dim=2).sum())
    return pairs / issued, pairs
tot = {1: 0.0, 2: 0.0, 3: 0.0, "u": 0.0}
for kmap, cin, cout in rows:
    if kmap.K != 27 or kmap.n_out < 1024 or not (cin >= 64 and cout >= 64): continue
    if id(kmap) not in seen:
        seen[id(kmap)] = {S: eff(kmap, S) for S in (1, 2, 3)}
        print(kmap.n_out, {S: round(v[0], 3) for S, v in seen[id(kmap)].items()})
    e = seen[id(kmap)]
    fl = e[1][1] * cin * cout
    tot["u"] += fl
    for S in (1, 2, 3): tot[S] += fl / e[S][0]
print({S: round(tot["u"] / tot[S], 3) for S in (1, 2, 3)}, "issued work vs S=1:", {S: round(tot[S] / tot[1], 3) for S in (2, 3)})
