#!/usr/bin/env python
"""cProfile of the launch side of one encode+decode (after warm-up): where the Python/ctypes time of the
frame pipeline's launch thread goes.  python tools/host_profile.py [exact|tf32] [radius]"""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcc_b200 as pcc

mode = sys.argv[1] if len(sys.argv) > 1 else "tf32"
radius = float(sys.argv[2]) if len(sys.argv) > 2 else 252.0
pcc.ops.set_default_mode(pcc.ops.MODE_TF32 if mode == "tf32" else pcc.ops.MODE_EXACT)
model = pcc.ColorModel(pcc.OURS_CONFIG)
pcc.init_synthetic(model, 0)
model.eval().to("cuda:0")
model.update()
x = pcc.synthetic.shell_frame(radius).to("cuda:0")
Q = pcc.qmap.q_map(x, 0.5, 0.5)


def frame():
    s, sh, k, c = model.compress(x, Q)
    return model.decompress(coordinates=c, strings=s, shape=sh, k=k)


for _ in range(2):
    frame()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
frame()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
st.sort_stats("cumulative").print_stats(30)
