"""One encode+decode of the configs[1] frame for ncu (run under `ncu --profile-from-start off`): a warm-up
round trip, then one round trip between cudaProfilerStart/Stop.

  python tools/profile_frame.py [--radius 252] [--mode exact|tf32] [--config ours|ablation]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcc_b200 as pcc

ap = argparse.ArgumentParser()
ap.add_argument("--radius", type=float, default=252.0)
ap.add_argument("--grid", type=int, default=1024)
ap.add_argument("--mode", default="exact")
ap.add_argument("--config", default="ours")
args = ap.parse_args()
dev = "cuda:0"
pcc.ops.set_default_mode(pcc.ops.MODE_TF32 if args.mode == "tf32" else pcc.ops.MODE_EXACT)
model = pcc.ColorModel(pcc.OURS_CONFIG if args.config == "ours" else pcc.ABLATION_CONFIG)
pcc.init_synthetic(model, 0)
model.eval().to(dev)
model.update()
x = pcc.synthetic.shell_frame(args.radius, grid=args.grid, seed=0)
qc, qf = pcc.synthetic.uniform_q_map(x, 0.5, 0.5)
xd, Q = x.to(dev), (qc.to(dev), qf.to(dev))


def roundtrip():
    s, sh, k, c = model.compress(xd, Q)
    return model.decompress(coordinates=c, strings=s, shape=sh, k=k)


roundtrip()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
out = roundtrip()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("decoded", tuple(out.shape), "launches", pcc.ops.launches())
