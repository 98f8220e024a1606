"""Tile efficiency of the output-stationary sparse conv: pairs / (128 * sum over tiles of active taps),
for the lexicographic row order vs rows sorted by (split) tap mask.  CPU/numpy analysis only."""
import sys
import numpy as np
sys.path.insert(0, ".")
import pcc_b200

def keys_of(c):
    return (c[:, 0].astype(np.int64) << 40) | (c[:, 1].astype(np.int64) << 20) | c[:, 2].astype(np.int64)

def masks(coords, ts):
    k = keys_of(coords); order = np.argsort(k); ks = k[order]
    m = np.zeros(len(coords), np.uint32); t = 0
    for dz in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                q = keys_of(coords + np.array([dx, dy, dz]) * ts)
                pos = np.searchsorted(ks, q); pos[pos >= len(ks)] = 0
                m |= ((ks[pos] == q).astype(np.uint32) << t); t += 1
    return m

def eff(m, splits):
    pairs = sum(int(((m >> t) & 1).sum()) for t in range(27))
    n = len(m); work = 0
    per = (27 + splits - 1) // splits
    for s in range(splits):
        sub = (m >> (s * per)) & ((1 << per) - 1)
        if splits > 1 or True:
            o = np.argsort(sub, kind="stable") if splits >= 1 and sort else np.arange(n)
        sub = sub[o]
        pad = (-n) % 128
        subp = np.concatenate([sub, np.zeros(pad, sub.dtype)]).reshape(-1, 128)
        tile_or = np.bitwise_or.reduce(subp, axis=1)
        work += 128 * sum(int(((tile_or >> t) & 1).sum()) for t in range(per))
    return pairs / work, pairs / n

x = pcc_b200.synthetic.shell_frame(float(sys.argv[1]) if len(sys.argv) > 1 else 252.0, seed=0).numpy()[:, :3].astype(np.int64)
def lex(c):
    return c[np.lexsort((c[:, 2], c[:, 1], c[:, 0]))]
levels = {"L0": (lex(x), 1)}
c = x
for l in (1, 2, 3):
    ts = 1 << l
    c = np.unique((c // ts) * ts, axis=0)
    levels[f"L{l}"] = (lex(c), ts)
# candidate sets
for l in (3, 2, 1):
    c, ts = levels[f"L{l}"]
    half = ts // 2
    offs = np.array([[dx, dy, dz] for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]) * half
    cand = np.unique((c[:, None, :] + offs[None]).reshape(-1, 3), axis=0)
    levels[f"cand{l-1}"] = (lex(cand), half)
for name, (c, ts) in levels.items():
    m = masks(c, ts)
    sort = False; e0, avg = eff(m, 1)
    sort = True
    print(f"{name}: n={len(c)} avg nbrs {avg:.2f}  lexicographic {e0:.3f}  sorted S=1 {eff(m,1)[0]:.3f}  S=2 {eff(m,2)[0]:.3f}  S=3 {eff(m,3)[0]:.3f}  S=4 {eff(m,4)[0]:.3f}")

def eff_seg(m, seg):
    pairs = sum(int(((m >> t) & 1).sum()) for t in range(27))
    n = len(m)
    blk = np.arange(n) // seg
    o = np.lexsort((m, blk))
    sub = m[o]
    pad = (-n) % 128
    subp = np.concatenate([sub, np.zeros(pad, sub.dtype)]).reshape(-1, 128)
    tile_or = np.bitwise_or.reduce(subp, axis=1)
    work = 128 * sum(int(((tile_or >> t) & 1).sum()) for t in range(27))
    return pairs / work
print("segmented mask sort (rows sorted by mask inside blocks of SEG lexicographic rows)")
for name, (c, ts) in levels.items():
    m = masks(c, ts)
    print(name, " ".join(f"seg{seg}: {eff_seg(m, seg):.3f}" for seg in (2048, 8192, 32768, 131072)))
