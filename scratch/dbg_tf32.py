import sys, numpy as np, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import pcc_b200 as pcc
from conftest import random_cloud
ops = pcc.ops; DEV="cuda:0"
def run(n, cin, cout, ext):
    rng = np.random.default_rng(1)
    c = random_cloud(n, ext, 3, 1, -2)
    m, _ = ops.canonical_map(torch.from_numpy(c).to(DEV))
    x = torch.randint(-4, 5, (m.n, cin), device=DEV).float()
    W = torch.randint(-2, 3, (27, cin, cout), device=DEV).float()
    km = m.kernel_map(m, 3)
    ref = ops.conv(x, km, W, None, mode=ops.MODE_EXACT)
    outs = [ops.conv(x, km, W, None, mode=ops.MODE_TF32) for _ in range(3)]
    torch.cuda.synchronize()
    bad = (outs[0] != ref)
    print(f"n={m.n} cin={cin} cout={cout}: bad elems {int(bad.sum())} / {bad.numel()}, deterministic {torch.equal(outs[0],outs[1]) and torch.equal(outs[1],outs[2])}")
    if bad.any():
        rows = bad.any(1).nonzero().flatten().cpu().numpy()
        tiles = np.unique(rows // 128)
        print("  bad tiles", tiles[:20], "of", (m.n+127)//128, " bad rows in first bad tile:", rows[rows//128==tiles[0]][:10] % 128, "count", (rows//128==tiles[0]).sum())
        cols = bad.any(0).nonzero().flatten().cpu().numpy()
        print("  bad cols", cols[:8], "...", cols[-4:], "count", len(cols))
        r0 = rows[0]; print("  sample row", r0, "ref", ref[r0,:6].tolist(), "got", outs[0][r0,:6].tolist())
        tm = km.tapmask.cpu().numpy().view(np.uint32)
        print("  popcount(tapmask) of bad tiles", [bin(int(tm[t])).count("1") for t in tiles[:10]], " of good tiles", [bin(int(tm[t])).count("1") for t in range(len(tm)) if t not in set(tiles)][:10])
for n, ext in [(100,12),(300,12),(1000,16),(5000,26)]:
    for cin, cout in [(64,64),(128,128),(128,64),(64,128),(96,128),(128,256)]:
        run(n, cin, cout, ext)
