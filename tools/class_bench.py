"""Roofline of every NON-dense kernel class of the path, on the maps of the configs[1]
frame: achieved = ALGORITHMIC bytes (the survey's per-unit figures x units) / CUDA-event time, against the
measured HBM copy bandwidth (MEASURED_PEAKS.json).  Used by bench.py (roofline.classes) and, with --profile, as
the program for one `ncu --set full --profile-from-start off` capture (every class runs exactly once inside
cudaProfilerStart/Stop, so the capture holds ~25 launches).

  python tools/class_bench.py [--radius 252] [--reps 5] [--profile] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _time(fn, reps, setup=None):
    """Median CUDA-event time (ms) of fn() over reps, after one warm-up; setup() runs untimed before each call."""
    ts = []
    for i in range(reps + 1):
        if setup is not None:
            setup()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i:
            ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def measure_classes(device="cuda:0", radius=252.0, grid=1024, reps=5, profile=False, hbm_gbs=None):
    """-> list of {"class", "kernels", "units", "bytes", "ms", "gbs", "frac"} (frac of hbm_gbs)."""
    import pcc_b200 as pcc
    ops = pcc.ops
    if hbm_gbs is None:
        try:
            hbm_gbs = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        except Exception:
            hbm_gbs = 6650.0  # B200_PROFILING.md fallback
    g = torch.Generator().manual_seed(0)
    x = pcc.synthetic.shell_frame(radius, grid=grid, seed=0)
    c0 = torch.cat([torch.zeros((x.shape[0], 1)), x[:, :3]], 1).int().to(device)
    m0, _ = ops.canonical_map(c0, 1)            # L0: 799,458 voxels
    m1 = m0.stride2()                           # L1: 249,574
    cand = m1.expand(3)                         # generative candidates of the last decoder stage (~4 M)
    n0, n1, nc = m0.n, m1.n, cand.n
    out = []
    todo = []  # everything is prepared first, so that a profiler range holds exactly one execution per class

    def add(name, kernels, units, nbytes, fn, setup=None):
        todo.append((name, kernels, units, nbytes, fn, setup))

    def fresh(keys, ts):
        return ops.CoordMap(keys, ts, True, 1)

    # ---- hash build: 16 B key + 8 B slot + 4 B value per coordinate
    for m, tag in ((m0, "L0 799k"), (cand, "candidates 4M")):
        add(f"hash build ({tag})", "hash_fill_kernel, hash_insert_kernel", m.n, 28.0 * m.n,
            lambda m=m: fresh(m.keys, m.ts).hash())
    # ---- kernel map: per output row 16 B + K x 8 B probes + 8 B per emitted pair (mask pass + order + write pass)
    for m, tag in ((m0, "L0 799k"), (cand, "candidates 4M")):
        m.hash()
        holder = {}

        def km(m=m):
            holder["k"] = fresh(m.keys, m.ts)
            holder["k"]._hash = m._hash
            return holder["k"].kernel_map(holder["k"], 3)
        kmap = km()
        pairs = int((kmap.nbr[:, :m.n] >= 0).sum().item())
        add(f"kernel map k=3 ({tag})", "kernel_map_kernel<3,1>, group_key_kernel, cub radix sort, tile_efficiency_kernel, "
            "kernel_map_kernel<3,0>, tile_rank_kernel", m.n, m.n * (16.0 + 27 * 8.0) + 8.0 * pairs, km)
    # ---- stride-2 map / generative map / canonical sort / pruning: 16 B in + 16 B out per coordinate
    add("stride-2 map (L0 -> L1)", "stride_keys_kernel, cub radix sort, head_flags, scan, compact_heads", n0,
        16.0 * n0 + 16.0 * n1, lambda: fresh(m0.keys, 1).stride2())
    m1.hash()

    def ex():
        t = fresh(m1.keys, 2)
        t._hash = m1._hash
        return t.expand(3)
    add("generative map (L1 -> candidates)", "expand_count_kernel, scan, expand_emit_kernel, cub radix sort", n1,
        16.0 * n1 + 16.0 * nc, ex)
    keys0 = m0.keys[torch.randperm(n0, generator=g).to(device)]
    add("canonical sort (799k keys + perm)", "iota_kernel, cub radix sort (8 passes of 64-bit keys)", n0,
        8 * (8 + 4) * 2.0 * n0, lambda: ops.sort_keys(keys0))
    rows = torch.sort(torch.randperm(nc, generator=g)[:n0]).values.to(torch.int32).to(device)
    add("prune keys (take 799k of 4M)", "take_keys_kernel", n0, 16.0 * n0 + 4.0 * n0, lambda: ops.take_keys(cand.keys, rows))
    # ---- top-k radix select: 4 B x M x <= 4 histogram passes + flags
    logits = torch.randn((nc, 1), generator=g).to(device)
    add("top-k select (799k of 4M)", "topk_init, topk_hist x4, topk_find x4, topk_eq/sel_flags, scan, topk_emit", nc,
        4.0 * nc * 4 + 1.0 * nc, lambda: ops.topk_rows(logits, n0))
    # ---- quantise + CDF index: 12 B read + 8 B write per element (the latent is small: 17,957 x 128)
    m3 = m1.stride2().stride2()
    n3 = m3.n
    y = torch.randn((n3, 128), generator=g).to(device)
    par = torch.rand((n3, 256), generator=g).to(device)
    table = pcc.compressai_compat.get_scale_table().to(device)
    add("quantise + CDF index (17,957 x 128)", "gc_quantize_kernel<0>", n3 * 128, 20.0 * n3 * 128,
        lambda: ops.gc_quantize_index(y, par, None, table, 0.11))
    # ---- gather rows (features_at_coordinates / pruning): 2 x 4 x C bytes per row
    f64 = torch.randn((nc, 64), generator=g).to(device)
    add("gather rows (799k x 64 of 4M)", "gather_rows_kernel<1>", n0, 2 * 4.0 * 64 * n0 + 4.0 * n0,
        lambda: ops.gather_rows(f64, rows))
    # ---- small-channel convolutions: 4 (pairs Cin + N_out Cout) + 8 pairs
    km0 = m0.kernel_map(m0, 3)
    p0 = int((km0.nbr[:, :n0] >= 0).sum().item())
    kmc = cand.kernel_map(cand, 3)
    pc = int((kmc.nbr[:, :nc] >= 0).sum().item())
    for cin, cout, m, kmap, pairs, tag, kern in (
            (4, 64, m0, km0, p0, "L0 799k", "conv_exact_kernel<64,1,1,4> (thin input)"),
            (2, 2, m0, km0, p0, "L0 799k", "conv_exact_kernel<16,0,1,4>"),
            (32, 3, m0, km0, p0, "L0 799k", "conv_head_kernel"),
            (64, 4, cand, kmc, pc, "candidates 4M", "conv_head_kernel")):
        f = torch.randn((m.n, cin), generator=g).to(device)
        w = (torch.randn((27, cin, cout), generator=g) * 0.1).to(device)
        b = torch.zeros((1, cout)).to(device)
        add(f"small-channel conv {cin}->{cout} ({tag})", kern, pairs, 4.0 * (pairs * cin + m.n * cout) + 8.0 * pairs,
            lambda f=f, kmap=kmap, w=w, b=b: ops.conv(f, kmap, w, b, mode=ops.MODE_EXACT))
    if profile:
        for _, _, _, _, fn, setup in todo:  # warm (allocator, lazy module loading) outside the range
            if setup is not None:
                setup()
            fn()
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStart()
    for name, kernels, units, nbytes, fn, setup in todo:
        if profile:
            if setup is not None:
                setup()
            fn()
            torch.cuda.synchronize()
            ms = float("nan")
        else:
            ms = _time(fn, reps, setup)
        gbs = nbytes / (ms * 1e-3) / 1e9 if ms == ms else float("nan")
        out.append({"class": name, "kernels": kernels, "units": units, "bytes": nbytes, "ms": ms, "gbs": gbs,
                    "frac": gbs / hbm_gbs if gbs == gbs else None})
    if profile:
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStop()
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--radius", type=float, default=252.0)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--profile", action="store_true")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    res = measure_classes(radius=a.radius, reps=a.reps, profile=a.profile)
    for r in res:
        print(f'{r["class"]:44s} {r["units"]:>10d} units  {r["bytes"] / 1e6:9.1f} MB  {r["ms"]:8.3f} ms  '
              f'{r["gbs"]:8.1f} GB/s  {(r["frac"] or 0) * 100:5.1f} % of HBM copy   [{r["kernels"]}]')
    if a.json:
        json.dump(res, open(a.json, "w"), indent=1)
