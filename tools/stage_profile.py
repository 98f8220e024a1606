#!/usr/bin/env python
"""Per-stage wall times (device-synchronised) of one encode+decode of the 800k frame.
  python tools/stage_profile.py [--mode exact|tf32] [--reps 3]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pcc_b200
from pcc_b200 import codec

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="exact")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--radius", type=float, default=252.0)
a = ap.parse_args()
dev = "cuda:0"
pcc_b200.ops.set_default_mode(pcc_b200.ops.MODE_TF32 if a.mode == "tf32" else pcc_b200.ops.MODE_EXACT)
model = pcc_b200.ColorModel(pcc_b200.OURS_CONFIG)
pcc_b200.init_synthetic(model, 0)
model.eval().to(dev); model.update()
x = pcc_b200.synthetic.shell_frame(a.radius, seed=0)
qc, qf = pcc_b200.synthetic.uniform_q_map(x, 0.5, 0.5)
xd, qd = x.to(dev), (qc.to(dev), qf.to(dev))
for _ in range(2):
    s, sh, k, c = model.compress(xd, qd); model.decompress(coordinates=c, strings=s, shape=sh, k=k)
torch.cuda.synchronize()
codec.STAGE_MS = {}
l0 = pcc_b200.ops.launches(); t0 = time.perf_counter()
for _ in range(a.reps):
    s, sh, k, c = model.compress(xd, qd); model.decompress(coordinates=c, strings=s, shape=sh, k=k)
torch.cuda.synchronize()
tot = (time.perf_counter() - t0) * 1e3 / a.reps
st = {k_: round(v / a.reps, 2) for k_, v in codec.STAGE_MS.items()}
codec.STAGE_MS = None
print(json.dumps({"mode": a.mode, "total_ms_with_stage_syncs": round(tot, 1), "sum_stages": round(sum(st.values()), 1),
                  "launches": (pcc_b200.ops.launches() - l0) // a.reps, "stages_ms": st}, indent=1))
