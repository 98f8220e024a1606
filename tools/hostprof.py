#!/usr/bin/env python
"""cProfile of the host side of one encode+decode (tf32 mode: shortest kernels, host most exposed)."""
import cProfile, pstats, os, sys, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, pcc_b200
mode = sys.argv[1] if len(sys.argv) > 1 else "tf32"
dev = "cuda:0"
pcc_b200.ops.set_default_mode(pcc_b200.ops.MODE_TF32 if mode == "tf32" else pcc_b200.ops.MODE_EXACT)
model = pcc_b200.ColorModel(pcc_b200.OURS_CONFIG); pcc_b200.init_synthetic(model, 0); model.eval().to(dev); model.update()
x = pcc_b200.synthetic.shell_frame(252.0, seed=0); qc, qf = pcc_b200.synthetic.uniform_q_map(x, 0.5, 0.5)
xd, qd = x.to(dev), (qc.to(dev), qf.to(dev))
def run():
    s, sh, k, c = model.compress(xd, qd); model.decompress(coordinates=c, strings=s, shape=sh, k=k)
for _ in range(3): run()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(3): run()
torch.cuda.synchronize(); pr.disable()
st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("tottime").print_stats(28); print(st.getvalue()[:6000])
