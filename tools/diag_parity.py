"""Diagnostic (B200 + host oracle): where does the CUDA path first differ from the oracle on a full frame?
  python tools/diag_parity.py [ours|ablation] [radius]
Walks h_s layer by layer on both sides after a normal compress()."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcc_b200 as pcc
import oracle
from oracle import prims
from oracle.prims import OT, conv, conv_transpose, leaky_relu, features_at
from pcc_b200.codec import apply_conv, ST, LEAKY

cfg_name = sys.argv[1] if len(sys.argv) > 1 else "ablation"
radius = float(sys.argv[2]) if len(sys.argv) > 2 else 252.0
cfg = pcc.ABLATION_CONFIG if cfg_name == "ablation" else pcc.OURS_CONFIG
dev = "cuda:0"
prims.set_num_threads(len(os.sched_getaffinity(0)))
model = pcc.ColorModel(cfg)
pcc.init_synthetic(model, 0)
model.eval()
params = pcc.export_params(model)
model.to(dev)
model.update()
x = pcc.synthetic.shell_frame(radius, seed=0)
qc, qf = pcc.synthetic.uniform_q_map(x, 0.5, 0.5)
strings, shape, k, coords = model.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
em = model.entropy_model
oc = oracle.OracleCodec(params, cfg)
o_strings, o_shape, o_k, o_coords = oc.compress(x.numpy(), qc.numpy(), qf.numpy())


def cmp(name, a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    if a.shape != b.shape:
        print(f"{name}: SHAPE {a.shape} vs {b.shape}")
        return
    av, bv = (a.view(np.uint32), b.view(np.uint32)) if a.dtype == np.float32 else (a, b)
    bad = np.argwhere(av != bv)
    print(f"{name}: shape {a.shape} mismatches {bad.shape[0]}")
    for idx in bad[:8]:
        t = tuple(idx)
        print("   ", t, repr(a[t]), repr(b[t]))


cmp("y", model.last["y"].F.cpu().numpy(), oc.trace["y"].F)
cmp("z_hat", em.last["z_hat"].F.cpu().numpy(), oc.trace["z_hat"].F)
P = em.last["params"].F.cpu().numpy()
cmp("scales", P[:, :128].T, oc.trace["scales"])
cmp("means", P[:, 128:].T, oc.trace["means"])
cmp("indexes", em.last["indexes"].cpu().numpy(), oc.trace["indexes"])
print("strings equal:", strings[0][0] == o_strings[0][0], strings[1][0] == o_strings[1][0])

# h_s layer by layer
z_hat = em.last["z_hat"]
zo = oc.trace["z_hat"]
y_map = model.last["y"].m
g = apply_conv(z_hat, em.h_s[0])
go = oc._conv(zo, "entropy_model.h_s.0", 3)
cmp("h_s.0", g.F.cpu().numpy(), go.F)
g = apply_conv(g, em.h_s[1], out_map=g.m.expand(2), transpose=True, act=LEAKY)
go = oc._convT(go, "entropy_model.h_s.1", 2)
go = go.with_feats(leaky_relu(go.F))
cmp("h_s.1 coords", g.m.coords.cpu().numpy(), go.C)
cmp("h_s.1", g.F.cpu().numpy(), go.F)
g = apply_conv(g, em.h_s[3])
go = oc._conv(go, "entropy_model.h_s.3", 3)
cmp("h_s.3", g.F.cpu().numpy(), go.F)
km = g.m.kernel_map(g.m, 3)
print("h_s.3 map n", g.m.n, "n_groups", km.n_groups)
g = apply_conv(g, em.h_s[4], out_map=g.m.expand(2), transpose=True, act=LEAKY)
go = oc._convT(go, "entropy_model.h_s.4", 2)
go = go.with_feats(leaky_relu(go.F))
cmp("h_s.4 coords", g.m.coords.cpu().numpy(), go.C)
cmp("h_s.4", g.F.cpu().numpy(), go.F)
full = apply_conv(g, em.h_s[6])
fo = oc._conv(go, "entropy_model.h_s.6", 3)
cmp("h_s.6 full", full.F.cpu().numpy(), fo.F)
km = g.m.kernel_map(g.m, 3)
print("h_s.6 full map n", g.m.n, "n_groups", km.n_groups)
at = apply_conv(g, em.h_s[6], out_map=y_map)
cmp("h_s.6 at y", at.F.cpu().numpy(), features_at(fo, oc.trace["y"].C))
km = g.m.kernel_map(y_map, 3)
print("h_s.6 at-y map n_out", y_map.n, "n_groups", km.n_groups)
