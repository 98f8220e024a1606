"""Surface (ragged) levels of the configs[1] frame: one 128 -> 128, k = 3 convolution per stride-2^level map, with the
27 offsets in 1 / 2 / 3 / 9 groups (each group = its own mask-sorted tile rows, chained launches).  Prints time,
TFLOP/s on real pairs and the share of the padded tile MACs that are useful.

  python tools/ragged_conv_bench.py [--levels 1 2 3] [--groups 1 2 3 9] [--cin 128] [--cout 128] [--reps 7]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcc_b200 as pcc

ap = argparse.ArgumentParser()
ap.add_argument("--levels", type=int, nargs="+", default=[1, 2, 3])
ap.add_argument("--groups", type=int, nargs="+", default=[1, 2, 3, 9])
ap.add_argument("--cin", type=int, default=128)
ap.add_argument("--cout", type=int, default=128)
ap.add_argument("--reps", type=int, default=7)
a = ap.parse_args()
dev = "cuda:0"
ops = pcc.ops
x = pcc.synthetic.shell_frame(252.0, seed=0)
c0 = torch.cat([torch.zeros((x.shape[0], 1)), x[:, :3]], 1).int().to(dev)
m, _ = ops.canonical_map(c0, 1)
peak = 148 * 128 * 2 * 1.965e9 / 1e12
lv = 0
gen = torch.Generator().manual_seed(0)
while lv <= max(a.levels):
    if lv in a.levels:
        f = torch.randn((m.n, a.cin), generator=gen).to(dev)
        w = (torch.randn((27, a.cin, a.cout), generator=gen) * 0.05).to(dev)
        b = torch.zeros((1, a.cout)).to(dev)
        ref = None
        for g in a.groups:
            km = m.kernel_map(m, 3, groups=g)
            pairs = int((km.nbr[:, :m.n] >= 0).sum().item())
            tiles = (m.n + 127) // 128
            tm = km.tapmask.view(km.n_groups, tiles)
            padded = 0
            for gi, (lo, hi) in enumerate(km.group_ranges()):
                bits = tm[gi].to(torch.int64) & 0xffffffff
                cnt = torch.zeros_like(bits)
                for k in range(lo, hi):
                    cnt += (bits >> k) & 1
                padded += int(cnt.sum().item()) * 128
            out = ops.conv(f, km, w, b, mode=ops.MODE_EXACT)
            if ref is None:
                ref = out
            same = bool(torch.equal(ref, out))
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
            fl = 2.0 * pairs * a.cin * a.cout
            print(f"level {lv} rows {m.n} pairs/row {pairs / m.n:.2f} groups {km.n_groups}: {ms:.3f} ms "
                  f"{fl / ms / 1e9:.2f} TFLOP/s = {fl / ms / 1e9 / peak:.3f}; useful MAC share {pairs / padded:.3f}; "
                  f"padded-MAC rate {2.0 * padded * a.cin * a.cout / ms / 1e9 / peak:.3f}; bit-identical {same}", flush=True)
    m = m.stride2()
    lv += 1
