"""The dominant launch alone: 128 -> 128 channels, k = 3, on the generative candidates of the stride-4 level of the
configs[1] frame (~1.1 M rows, the shape of the decoder's second up-sampling stage).  Prints TFLOP/s and the
fraction of the FP32 FMA peak; with --profile runs exactly one launch inside cudaProfilerStart/Stop for
`ncu --set full --import-source on --profile-from-start off -k regex:conv_exact`.

  python tools/dense_conv_bench.py [--cin 128] [--cout 128] [--level 2] [--reps 5] [--profile]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcc_b200 as pcc

ap = argparse.ArgumentParser()
ap.add_argument("--cin", type=int, default=128)
ap.add_argument("--cout", type=int, default=128)
ap.add_argument("--level", type=int, default=2, help="stride-2^level map whose generative candidates are used")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--profile", action="store_true")
a = ap.parse_args()
dev = "cuda:0"
ops = pcc.ops
x = pcc.synthetic.shell_frame(252.0, seed=0)
c0 = torch.cat([torch.zeros((x.shape[0], 1)), x[:, :3]], 1).int().to(dev)
m, _ = ops.canonical_map(c0, 1)
for _ in range(a.level):
    m = m.stride2()
cand = m.expand(3)
km = cand.kernel_map(cand, 3)
pairs = int((km.nbr[:, :cand.n] >= 0).sum().item())
g = torch.Generator().manual_seed(0)
f = torch.randn((cand.n, a.cin), generator=g).to(dev)
w = (torch.randn((27, a.cin, a.cout), generator=g) * 0.05).to(dev)
b = torch.zeros((1, a.cout)).to(dev)
out = ops.conv(f, km, w, b, mode=ops.MODE_EXACT)
torch.cuda.synchronize()
flops = 2.0 * pairs * a.cin * a.cout
alg_bytes = 4.0 * (cand.n * a.cin + cand.n * a.cout) + 4.0 * 27 * cand.n + 4.0 * 27 * a.cin * a.cout
if a.profile:
    torch.cuda.cudart().cudaProfilerStart()
    ops.conv(f, km, w, b, mode=ops.MODE_EXACT)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print(f"rows {cand.n} pairs {pairs} groups {km.n_groups} algorithmic bytes {alg_bytes:.0f}")
else:
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.conv(f, km, w, b, mode=ops.MODE_EXACT)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    peak = 148 * 128 * 2 * 1.965e9 / 1e12
    print(f"rows {cand.n} pairs {pairs} groups {km.n_groups}: {ms:.3f} ms, {flops / ms / 1e9:.2f} TFLOP/s = "
          f"{flops / ms / 1e9 / peak:.3f} of {peak:.2f}; algorithmic bytes {alg_bytes / 1e6:.1f} MB")
