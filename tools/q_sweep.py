"""configs[1]'s quality sweep: encode+decode throughput of the 799,458-point frame for (q_g, q_a) in {0, 0.5, 1}^2
(the reference sweeps 21 x 21 values, evaluate.py:71-72,95-97), same pipeline as bench.py, device-resident inputs.
  python tools/q_sweep.py [--frames 8] [--json profiles/r02_q_sweep.json]"""
import argparse
import hashlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcc_b200 as pcc

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=8)
ap.add_argument("--depth", type=int, default=4)
ap.add_argument("--json", default=None)
a = ap.parse_args()
dev = torch.device("cuda:0")
model = pcc.ColorModel(pcc.OURS_CONFIG)
pcc.init_synthetic(model, 0)
model.eval().to(dev)
model.update()
x = pcc.synthetic.shell_frame(252.0, seed=0)
xd = x.to(dev)
pipe = pcc.pipeline.FramePipeline(model, depth=a.depth, device=dev, scheduler="threads")
rows = []
for qg in (0.0, 0.5, 1.0):
    for qa in (0.0, 0.5, 1.0):
        qc, qf = pcc.synthetic.uniform_q_map(x, qg, qa)
        frame = (xd, (qc.to(dev), qf.to(dev)))
        pipe.roundtrip_many([frame] * a.depth)  # warm-up
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res = pipe.roundtrip_many([frame] * a.frames)
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1) * 1e-3
        s, sh, k, c = model.compress(*frame)
        rows.append({"q_g": qg, "q_a": qa, "points_per_s": x.shape[0] * a.frames / t, "ms_per_frame": t / a.frames * 1e3,
                     "y_bytes": len(s[0][0]), "z_bytes": len(s[1][0]), "bpp": 8.0 * (len(s[0][0]) + len(s[1][0])) / x.shape[0],
                     "decoded_points": int(res[0][0].shape[0]),
                     "strings_sha1": hashlib.sha1(s[0][0] + s[1][0]).hexdigest()})
        print(rows[-1])
pipe.close()
vals = [r["points_per_s"] for r in rows]
out = {"workload": "Ours.yaml random-init, 799,458-point 10-bit shell frame, exact mode, in-memory variant",
       "frames_per_point": a.frames, "frames_in_flight": a.depth, "sweep": rows,
       "spread": (max(vals) - min(vals)) / (sum(vals) / len(vals))}
print("spread of throughput over the sweep: %.2f %%" % (100 * out["spread"]))
if a.json:
    json.dump(out, open(a.json, "w"), indent=1)
