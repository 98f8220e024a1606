#!/usr/bin/env python
"""Hunts for intermittent slow steps: per-rep stage times, flags stages > 1.5x their median."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import pcc_b200
from pcc_b200 import codec
mode = sys.argv[1] if len(sys.argv) > 1 else "exact"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
host = len(sys.argv) > 3 and sys.argv[3] == "host"
dev = "cuda:0"
pcc_b200.ops.set_default_mode(pcc_b200.ops.MODE_TF32 if mode == "tf32" else pcc_b200.ops.MODE_EXACT)
model = pcc_b200.ColorModel(pcc_b200.OURS_CONFIG); pcc_b200.init_synthetic(model, 0); model.eval().to(dev); model.update()
x = pcc_b200.synthetic.shell_frame(252.0, seed=0); qc, qf = pcc_b200.synthetic.uniform_q_map(x, 0.5, 0.5)
if host:
    xd, qd = x.pin_memory(), (qc.pin_memory(), qf.pin_memory())
else:
    xd, qd = x.to(dev), (qc.to(dev), qf.to(dev))
for _ in range(3):
    s, sh, k, c = model.compress(xd, qd); model.decompress(coordinates=c, strings=s, shape=sh, k=k)
import gc; gc.collect(); gc.disable()
rows = []
for r in range(reps):
    codec.STAGE_MS = {}
    torch.cuda.synchronize(); t0 = time.perf_counter()
    s, sh, k, c = model.compress(xd, qd)
    if host: c = c.cpu()
    o = model.decompress(coordinates=c, strings=s, shape=sh, k=k)
    if host: o = o.cpu()
    torch.cuda.synchronize(); tot = (time.perf_counter() - t0) * 1e3
    d = dict(codec.STAGE_MS); d["TOTAL"] = tot; rows.append(d)
codec.STAGE_MS = None
keys = list(rows[0].keys())
med = {k_: float(np.median([r[k_] for r in rows])) for k_ in keys}
print("median", {k_: round(v, 1) for k_, v in med.items()})
for i, r in enumerate(rows):
    bad = {k_: round(r[k_], 1) for k_ in keys if r[k_] > 1.5 * med[k_] + 2}
    print(i, round(r["TOTAL"], 1), bad)
print("memory reserved GB", torch.cuda.memory_reserved() / 2**30, "num_alloc_retries", torch.cuda.memory_stats().get("num_alloc_retries"), "segments", torch.cuda.memory_stats().get("segment.all.allocated"))
