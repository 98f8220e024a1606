#!/usr/bin/env python
"""Headline benchmark: encode+decode points/sec on a synthetic ~800k-point 10-bit frame
(BASELINE.json metric; configs[1]: Ours.yaml architecture, random-init weights, shell radius 252
-> 799,458 voxels, RGB, uniform quality map).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # CPU restatement of the reference path

A step = ColorModel.compress + ColorModel.decompress of one frame (in-memory variant, path=None).
N > 1 (torchrun): frames are independent, so each rank codes its own frame of the synthetic
sequence (configs[3]) with no data-path collective: weak scaling, value = sum of points / max time.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "encode+decode points/sec, 800k-pt 10-bit frame"
UNIT = "points/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default=os.environ.get("PCC_BENCH_MODE", "exact"), choices=["exact", "tf32"])
    ap.add_argument("--radius", type=float, default=252.0, help="shell radius (252 -> 799,458 voxels)")
    ap.add_argument("--grid", type=int, default=1024, help="voxel grid edge (1024 = 10 bit; 4096 = 12 bit, configs[4])")
    ap.add_argument("--cpu-radius", type=float, default=89.0, help="shell radius of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--config", default="ours", choices=["ours", "ablation"])
    ap.add_argument("--q", type=float, nargs=2, default=[0.5, 0.5], metavar=("Q_G", "Q_A"))
    ap.add_argument("--dump-layers", default=None, help="write the per-layer conv timing table here")
    ap.add_argument("--frames-per-step", type=int, default=12,
                    help="frames coded per step per GPU (frame pipeline; 1 = one frame at a time)")
    ap.add_argument("--depth", type=int, default=0,
                    help="frames in flight per GPU; 0 = 6 (measured on the 16-core box, 799,458-point frame: depth 4 / 6 / 8 = "
                         "4.90 / 5.13 / - M points/s exact, 15.7 / 17.9 / 15.3 M tf32)")
    ap.add_argument("--scheduler", default="auto", choices=["auto", "threads", "cooperative"],
                    help="frame pipeline: one launch thread per in-flight frame, or one thread resuming all frames; "
                         "auto = threads when this rank's share of the host cores is at least 4")
    ap.add_argument("--no-tf32-pass", action="store_true",
                    help="skip the secondary tensor-core (tf32) measurement added to the exact-mode line")
    ap.add_argument("--no-scene-pass", action="store_true",
                    help="N > 1: skip the secondary configs[4] record (one 5 M-point scene over the same ranks)")
    ap.add_argument("--scene", action="store_true",
                    help="configs[4]: ONE 12-bit ~5 M-point scene cut into x-slabs across the GPUs (halo exchange "
                         "per layer, strong scaling) instead of the frame-parallel headline workload")
    ap.add_argument("--scene-radius", type=float, default=630.0)
    ap.add_argument("--scene-grid", type=int, default=4096)
    ap.add_argument("--verify", action="store_true",
                    help="--scene: also code the whole scene on rank 0 alone and compare strings and decoded cloud")
    return ap.parse_args()


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi sampling DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, n_gpus=1):
        # every GPU of the job is sampled (rank r uses GPU r of the node): the step time is the max over ranks, so
        # one throttled GPU is enough to slow the job
        self.gpus = list(range(n_gpus)) if n_gpus > 1 else [gpu_index]
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", ",".join(str(g) for g in self.gpus)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = {}, [], {}, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [v.strip() for v in line.split(",")]
            if len(f) < 9:
                continue
            try:
                g = int(f[0])
                sm.setdefault(g, []).append(float(f[1]))
                smax.append(float(f[2]))
                power.setdefault(g, []).append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        med = {g: sorted(v)[len(v) // 2] for g, v in sm.items()}
        out = {"sm_mhz": min(med.values()) if med else None, "sm_max_mhz": max(smax) if smax else None,
               "samples": sum(len(v) for v in sm.values()), "reasons": sorted(reasons)}
        if len(med) > 1:  # sm_mhz = the slowest GPU's median under load
            out["per_gpu_sm_mhz"] = [med[g] for g in sorted(med)]
            out["per_gpu_power_w"] = [round(sorted(power[g])[len(power[g]) // 2], 1) for g in sorted(power)]
        return out


# ------------------------------------------------------------------------------ workload
def make_model(cfg_name, device=None):
    import pcc_b200
    cfg = pcc_b200.OURS_CONFIG if cfg_name == "ours" else pcc_b200.ABLATION_CONFIG
    model = pcc_b200.ColorModel(cfg)
    pcc_b200.init_synthetic(model, 0)
    model.eval()
    if device is not None:
        model.to(device)
        model.update()
    return model, cfg


def frame_for(rank, step, world, radius, grid=1024):
    """Rank-sharded synthetic sequence: rank r codes frames r, r + world, ... (frame-parallel)."""
    import pcc_b200
    if world == 1:
        return pcc_b200.synthetic.shell_frame(radius, grid=grid, seed=0)
    t = pcc_b200.frames.frames_of_rank((step + 1) * world, rank, world)[step]  # = step * world + rank
    import numpy as np
    r = radius + 4.0 * np.sin(2.0 * np.pi * t / 300.0)
    return pcc_b200.synthetic.shell_frame(r, grid=grid, seed=t)


# ------------------------------------------------------------------------------ reference arm
def workload_name(args, n_frame):
    """The workload string both arms carry in ``config`` (same frame, same quality, same variant)."""
    return (f"Ours.yaml random-init ({args.config}), encode+decode of one synthetic {(args.grid - 1).bit_length()}-bit "
            f"voxelised shell frame, radius {args.radius:g} in a {args.grid}^3 grid -> {n_frame} points, RGB, "
            f"uniform q=({args.q[0]:g},{args.q[1]:g}); in-memory variant (path=None)")


def host_threads():
    """Every host core this process may run on.  torchrun exports OMP_NUM_THREADS=1; the CPU arm must not
    inherit that (rank 0 alone runs it, the other ranks exit), so the count is set explicitly."""
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def oracle_roundtrip(oc, x, qc, qf):
    t0 = time.perf_counter()
    s, sh, k, c = oc.compress(x, qc, qf)
    out = oc.decompress(c, s, sh, k)
    return time.perf_counter() - t0, (s, sh, k, c, out)


REFERENCE_BUDGET_S = 150.0  # timed steps stop once this much CPU time has been spent


def run_reference(args, rank):
    """The reference's own CPU implementation cannot be installed (MinkowskiEngine / CompressAI
    absent), so this arm times the CPU restatement (oracle/, kind "port") with all host threads on
    the SAME frame as the CUDA arm (configs[1]: 799,458 points).  A step = encode+decode of that frame
    (~25 s on the box); warm-up steps run on the ~100k-point frame of configs[0] (they only have to
    touch the code, the OpenMP pool and the entropy tables), and the timed steps stop after
    REFERENCE_BUDGET_S so that the run stays within a few minutes."""
    if rank != 0:
        return
    import pcc_b200
    from oracle import OracleCodec, prims
    prims.build()
    cores = host_threads()
    prims.set_num_threads(cores)
    model, cfg = make_model(args.config)
    oc = OracleCodec(pcc_b200.export_params(model), cfg)
    x = pcc_b200.synthetic.shell_frame(args.radius, grid=args.grid, seed=0)
    qc, qf = pcc_b200.synthetic.uniform_q_map(x, *args.q)
    xs = pcc_b200.synthetic.shell_frame(args.cpu_radius, seed=0)
    qcs, qfs = pcc_b200.synthetic.uniform_q_map(xs, *args.q)
    small = []
    for _ in range(max(0, args.warmup)):
        small.append(oracle_roundtrip(oc, xs.numpy(), qcs.numpy(), qfs.numpy())[0])
    want = max(1, args.steps)
    times = []
    while len(times) < want and (not times or sum(times) + times[-1] <= REFERENCE_BUDGET_S):
        times.append(oracle_roundtrip(oc, x.numpy(), qc.numpy(), qf.numpy())[0])
    steps = len(times)
    dt = sum(times) / steps
    value = x.shape[0] / dt
    sample = (f"{steps} timed step(s) of the full configs[1] frame ({x.shape[0]} points, {dt:.1f} s each; {want} asked, "
              f"bounded by {REFERENCE_BUDGET_S:g} s), {len(small)} warm-up step(s) on the configs[0] frame "
              f"({xs.shape[0]} points)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": max(0, args.warmup), "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, x.shape[0])},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if small:
        line["cpu_baseline"]["value_100k_frame"] = xs.shape[0] / min(small)
    print(json.dumps(line))


# ------------------------------------------------------------------------------ our arm
def conv_flops(r):
    return 2.0 * r["pairs"] * r["cin"] * r["cout"]


def conv_bytes(r):
    """Algorithmic bytes of one sparse conv: features in/out once, kernel map, weights."""
    return 4.0 * (r["n_in"] * r["cin"] + r["n_out"] * r["cout"]) + 4.0 * r["K"] * r["n_out"] \
        + 4.0 * r["K"] * r["cin"] * r["cout"]


def roofline_pass(model, x_dev, q_dev, peaks, clocks, dump=None):
    """One extra (untimed) step with CUDA events around every pcc_conv_fwd launch on the launching
    stream; the dominant kernel is the 128-wide sparse-conv kernel (all cin,cout >= 64 layers)."""
    import pcc_b200
    ops = pcc_b200.ops
    ops.start_conv_profile()
    s, sh, k, c = model.compress(x_dev, q_dev)
    model.decompress(coordinates=c, strings=s, shape=sh, k=k)
    recs = ops.stop_conv_profile()
    if not recs:
        return None, []
    big = [r for r in recs if r["cin"] >= 64 and r["cout"] >= 64]
    dom = big if big else recs
    t_ms = sum(r["ms"] for r in dom)
    fl = sum(conv_flops(r) for r in dom)
    tf32 = any(r["mode"] == 1 for r in dom)
    achieved = fl / (t_ms * 1e-3) / 1e12
    if tf32:
        peak = peaks.get("bf16_tflops", 1590.0) / 2.0
        bound, src = "tensor", "tf32 dense = measured bf16 burst / 2 (MEASURED_PEAKS.json)" \
            if "bf16_tflops" in peaks else "fallback 1590/2"
    else:
        mhz = clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz", 1965.0)
        peak = 148 * 128 * 2 * mhz * 1e6 / 1e12
        bound, src = "fp32_fma", f"148 SM x 128 FMA x 2 x {mhz:g} MHz (no measured FP32 peak exists)"
    roof = {"bound": bound, "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            "traffic": None,
            "kernel": "conv_tf32_kernel" if tf32 else "conv_exact_tma_kernel<128,8> / <64,8> (every launch with Cin, Cout >= 64)",
            "launches": len(dom), "avg_launch_ms": t_ms / len(dom), "peak_source": src,
            "share_of_conv_time": t_ms / sum(r["ms"] for r in recs),
            "hbm_gbs_algorithmic": sum(conv_bytes(r) for r in dom) / (t_ms * 1e-3) / 1e9}
    try:  # DRAM traffic of one captured launch (ncu --set full), committed under profiles/
        prof = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles")
        name = "r02_traffic.json" if os.path.exists(os.path.join(prof, "r02_traffic.json")) else "r01_traffic.json"
        with open(os.path.join(prof, name)) as f:
            tr = json.load(f)["tf32" if tf32 else "exact"]
        roof["traffic"] = tr["traffic_bytes"]
        roof["traffic_detail"] = {k: tr[k] for k in ("launch", "algorithmic_bytes", "source")}
    except (OSError, KeyError, ValueError):
        pass
    if dump:
        with open(dump, "w") as f:
            f.write("cin,cout,K,n_in,n_out,pairs,mode,ms,tflops,alg_GBs\n")
            for r in recs:
                f.write(f'{r["cin"]},{r["cout"]},{r["K"]},{r["n_in"]},{r["n_out"]},{r["pairs"]},{r["mode"]},'
                        f'{r["ms"]:.4f},{conv_flops(r) / (r["ms"] * 1e-3) / 1e12:.3f},'
                        f'{conv_bytes(r) / (r["ms"] * 1e-3) / 1e9:.1f}\n')
    return roof, recs


def tf32_pass(args, model, dev, frame_dev, B, peaks, clocks):
    """Secondary measurement in the tensor-core mode (tcgen05 kind::tf32 for every layer with
    Cin, Cout multiples of 32), same workload and pipeline; N = 1 only.  Not the parity mode: latents
    agree with the exact mode to the tolerance stated in tests/test_gpu_tf32.py."""
    import torch
    import pcc_b200
    ops = pcc_b200.ops
    ops.set_default_mode(ops.MODE_TF32)
    try:
        pipe = pcc_b200.pipeline.FramePipeline(model, depth=min(args.depth, B), device=dev, scheduler=args.scheduler)
        import gc
        steps, warm = 3, 3
        for _ in range(warm):
            pipe.roundtrip_many([frame_dev] * B)
        torch.cuda.synchronize()
        gc.collect()
        gc.disable()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        npts = 0
        wall = []
        for _ in range(steps):
            tw = time.perf_counter()
            npts += sum(o.shape[0] for o, _, _ in pipe.roundtrip_many([frame_dev] * B))
            wall.append(round((time.perf_counter() - tw) * 1e3, 2))
        ev1.record()
        torch.cuda.synchronize()
        gc.enable()
        t = ev0.elapsed_time(ev1) * 1e-3
        roof, _ = roofline_pass(model, *frame_dev, peaks, clocks, None)
        return {"value": npts / t, "unit": UNIT, "ms_per_frame": t / (steps * B) * 1e3, "steps": steps,
                "warmup": warm, "dtype": "tf32", "roofline": roof, "step_host_ms": wall,
                "note": "tcgen05.mma kind::tf32 path, same frames and pipeline; tolerance mode, not bit-exact"}
    finally:
        ops.set_default_mode(ops.MODE_EXACT)


def cpu_baseline(args, cfg_name, gpu_result=None):
    """Oracle (kind "port") timed on all host cores of the box (rank 0, N=1 only): one warmed-up
    encode+decode of the SAME 799,458-point frame the CUDA arm codes (the ~100k configs[0] frame is the
    warm-up and a second figure).  Its strings, latent coordinates and decoded cloud are compared with
    the CUDA arm's (gpu_result) -> "verified_bit_exact_vs_oracle"."""
    import numpy as np
    import pcc_b200
    from oracle import OracleCodec, prims
    prims.build()
    cores = host_threads()
    prims.set_num_threads(cores)
    model, cfg = make_model(cfg_name)
    oc = OracleCodec(pcc_b200.export_params(model), cfg)
    xs = pcc_b200.synthetic.shell_frame(args.cpu_radius, seed=0)
    qcs, qfs = pcc_b200.synthetic.uniform_q_map(xs, *args.q)
    oracle_roundtrip(oc, xs.numpy()[:2000], qcs.numpy()[:2000], qfs.numpy()[:2000])  # touch code, pool, tables
    dt_s, _ = oracle_roundtrip(oc, xs.numpy(), qcs.numpy(), qfs.numpy())
    x = pcc_b200.synthetic.shell_frame(args.radius, grid=args.grid, seed=0)
    qc, qf = pcc_b200.synthetic.uniform_q_map(x, *args.q)
    dt, (s, sh, k, c, out) = oracle_roundtrip(oc, x.numpy(), qc.numpy(), qf.numpy())
    base = {"value": x.shape[0] / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"one warmed-up encode+decode of the full configs[1] frame ({x.shape[0]} points, {dt:.1f} s); "
                      f"warm-up = the configs[0] frame ({xs.shape[0]} points, {dt_s:.1f} s)",
            "value_100k_frame": xs.shape[0] / dt_s}
    verified = None
    if gpu_result is not None:
        gs, gsh, gk, gc_, gout = gpu_result
        verified = bool(gs[0][0] == s[0][0] and gs[1][0] == s[1][0] and list(gsh) == list(sh) and gk == k
                        and np.array_equal(gc_, c) and gout.shape == out.shape
                        and np.array_equal(gout.view(np.uint32), out.view(np.uint32)))
    return base, verified


def run_scene(args, rank, world, local_rank):
    """BASELINE.json configs[4]: one large scene, spatially partitioned (pcc_b200.slab / .scene)."""
    import torch
    import pcc_b200
    ops = pcc_b200.ops
    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    ops.set_default_mode(ops.MODE_TF32 if args.mode == "tf32" else ops.MODE_EXACT)
    model, cfg = make_model(args.config, dev)
    line = measure_scene(args, rank, world, local_rank, model, args.steps, args.warmup, args.verify)
    if rank == 0:
        print(json.dumps(line))


def measure_scene(args, rank, world, local_rank, model, steps, warmup, verify):
    """One scene, x-slabs over the ranks (every rank calls this) -> the JSON line (rank 0) or None."""
    import hashlib
    import torch
    import torch.distributed as dist
    import pcc_b200
    ops, scene, slab_m = pcc_b200.ops, pcc_b200.scene, pcc_b200.slab

    dev = torch.device(f"cuda:{local_rank}")
    x = pcc_b200.synthetic.shell_frame(args.scene_radius, grid=args.scene_grid, seed=0)  # same on every rank
    Q = pcc_b200.synthetic.uniform_q_map(x, *args.q)
    n_scene = x.shape[0]
    comm = slab_m.DistComm() if world > 1 else slab_m.ThreadComm.make(1)[0]
    bounds = scene.balanced_bounds(x, world)
    xs, (qc, qf) = scene.split_scene(x, Q, bounds)[rank]
    sl = slab_m.Slab(comm, bounds)
    host = (xs.pin_memory(), qc.pin_memory(), qf.pin_memory())
    resident = (xs.to(dev), (qc.to(dev), qf.to(dev)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(host_io):
        if host_io:
            xd, Qd = host[0].to(dev, non_blocking=True), (host[1].to(dev, non_blocking=True), host[2].to(dev, non_blocking=True))
        else:
            xd, Qd = resident
        s, sh, k, c = scene.compress_scene(model, xd, Qd, sl)
        out = scene.decompress_scene(model, c, s, sh, k, sl, gather=False)
        if host_io:
            out = out.cpu()
        return s, c, out

    import gc
    res = {}
    for name, host_io in (("dev", False), ("e2e", True)):
        for _ in range(warmup if name == "dev" else 1):
            step(host_io)
        gc.collect()
        gc.disable()
        barrier()
        sampler = ClockSampler(local_rank, world)
        if rank == 0 and name == "dev":
            sampler.start()
        l0 = ops.launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for _ in range(steps):
            s, c, out = step(host_io)
        e1.record()
        barrier()
        gc.enable()
        t = max(e0.elapsed_time(e1) * 1e-3, (time.perf_counter() - t0) if host_io else 0.0)
        res[name] = {"t": t, "launches": ops.launches() - l0}
        if rank == 0 and name == "dev":
            res["clocks"] = sampler.stop()
    (t_dev, t_e2e), _ = pcc_b200.frames.reduce_throughput([res["dev"]["t"], res["e2e"]["t"]], 0, dev)
    own_pts = torch.tensor([xs.shape[0], out.shape[0]], dtype=torch.int64, device=dev)
    if world > 1:
        allp = [torch.empty_like(own_pts) for _ in range(world)]
        dist.all_gather(allp, own_pts)
        per_rank = [p.tolist() for p in allp]
    else:
        per_rank = [own_pts.tolist()]
    verified = None
    if verify:
        s2, sh2, k2, c2 = scene.compress_scene(model, *resident, sl)
        full = scene.decompress_scene(model, c2, s2, sh2, k2, sl, gather=True)
        if rank == 0:
            xa, Qa = x.to(dev), (Q[0].to(dev), Q[1].to(dev))
            rs, rsh, rk, rc = model.compress(xa, Qa)
            ref = model.decompress(coordinates=rc, strings=rs, shape=rsh, k=rk)
            verified = bool(rs[0][0] == s2[0][0] and rs[1][0] == s2[1][0] and rsh == sh2 and rk == k2
                            and torch.equal(rc, c2) and torch.equal(ref.view(torch.int32), full.view(torch.int32)))
    if rank != 0:
        return None
    h2d = (host[0].numel() + host[1].numel() + host[2].numel()) * 4
    d2h = out.numel() * 4
    line = {
        "metric": "encode+decode points/sec, 5M-pt 12-bit scene (configs[4])", "value": n_scene * steps / t_dev,
        "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": t_dev / steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if args.mode == "exact" else "tf32", "data": "synthetic",
        "config": {"workload": f"Ours.yaml random-init, encode+decode of ONE synthetic {args.scene_grid}^3-grid shell scene, "
                               f"radius {args.scene_radius:g} -> {n_scene} points, RGB, uniform q=({args.q[0]:g},{args.q[1]:g})",
                   "mode": args.mode, "parallelism": f"{world} x-slab(s), boundaries {bounds}, one-voxel halo ring per level, "
                   "neighbour send/recv per conv layer (NCCL), scene-wide top-k, latent gathered for the single rANS stream",
                   "points_in_out_per_rank": per_rank,
                   "l2": "working set exceeds the 126 MB L2; no flush", "includes_cpu_range_coder": True},
        "e2e": {"value": n_scene * steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": t_e2e / steps * 1e3, "note": "per rank: its slab H2D from pinned memory, its decoded slab D2H"},
        "gpu_launches": res["dev"]["launches"], "clocks": res.get("clocks", {}),
        "string_sha1": hashlib.sha1(s[0][0] + s[1][0]).hexdigest(),
    }
    if verified is not None:
        line["verified_bit_exact_vs_one_gpu"] = verified
    return line


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import pcc_b200
    from pcc_b200 import compressai_compat as cc
    ops = pcc_b200.ops

    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    ops.set_default_mode(ops.MODE_TF32 if args.mode == "tf32" else ops.MODE_EXACT)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    per_rank = cores // max(1, world)
    if args.depth <= 0:
        args.depth = 6
    if args.scheduler == "auto":
        # one launch thread per in-flight frame needs a few cores of its own (the threads wait on CUDA events and
        # on the coder); measured on ONE GPU pinned to 4 cores (taskset -c 0-3, 799,458-point frame, exact):
        # cooperative depth 4 / 6 / 8 = 4.72 / 4.84 / 4.83 M points/s, threads depth 3 / 4 / 6 = 4.80 / 4.97 / 5.05 M
        args.scheduler = "threads" if per_rank >= 4 else "cooperative"
    model, cfg = make_model(args.config, dev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass

    total = args.warmup + args.steps
    B = max(1, args.frames_per_step)
    # world == 1: the configs[1] frame, B copies per step; world > 1: rank r codes frames r, r + W, ...
    # of the synthetic sequence (configs[3]), B distinct frames per step
    frames = [frame_for(rank, s, world, args.radius, args.grid) for s in range(total * B if world > 1 else 1)]
    get = (lambda s: frames[s]) if world > 1 else (lambda s: frames[0])
    qmaps = {}
    pipe = pcc_b200.pipeline.FramePipeline(model, depth=min(args.depth, B), device=dev, scheduler=args.scheduler)

    def host_inputs(s):
        x = get(s)
        if id(x) not in qmaps:
            qc, qf = pcc_b200.synthetic.uniform_q_map(x, *args.q)
            qmaps[id(x)] = (x.pin_memory(), qc.pin_memory(), qf.pin_memory())
        return qmaps[id(x)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(s):
        """B frames already resident in HBM through the frame pipeline -> points coded."""
        res = pipe.roundtrip_many([dev_inputs[s * B + j] for j in range(B)])
        return sum(o.shape[0] for o, _, _ in res)

    def step_e2e(s):
        """B frames from pinned host memory; latent coordinates and decoded clouds back on the host."""
        ins = [host_inputs(s * B + j) for j in range(B)]
        res = pipe.roundtrip_many([(xh, (qch, qfh)) for xh, qch, qfh in ins], host_io=True)
        h2d = sum(xh.numel() * 4 + qch.numel() * 4 + qfh.numel() * 4 for xh, qch, qfh in ins) \
            + sum(c.numel() * 4 for _, _, c in res)
        d2h = sum(o.numel() * 4 + c.numel() * 4 for o, _, c in res)
        return h2d, d2h

    # ---- device-resident timing (value)
    dev_inputs = []
    for s in range(total * B):
        xh, qch, qfh = host_inputs(s)
        if world > 1 or s == 0:
            dev_inputs.append((xh.to(dev), (qch.to(dev), qfh.to(dev))))
        else:
            dev_inputs.append(dev_inputs[0])
    for s in range(args.warmup):
        step_device(s)
    import gc
    gc.collect()
    gc.disable()  # no collector pauses inside the timed regions (re-enabled after the e2e loop)
    barrier()
    sampler = ClockSampler(local_rank, world)
    if rank == 0:
        sampler.start()
    cc.reset_coder_stats()
    l0 = ops.launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    npts = 0
    step_wall, pipe_stats = [], []
    for s in range(args.warmup, total):
        tw = time.perf_counter()
        npts += step_device(s)
        step_wall.append(round((time.perf_counter() - tw) * 1e3, 2))
        pipe_stats.append({k: (round(v, 1) if isinstance(v, float) else v) for k, v in pipe.stats.items()})
    ev1.record()  # the pipeline made this stream wait for every frame's stream
    barrier()
    t_dev = ev0.elapsed_time(ev1) * 1e-3
    launches = ops.launches() - l0
    coder = dict(cc.coder_stats)
    clocks = sampler.stop() if rank == 0 else {}

    # ---- end-to-end timing through the public API with host buffers (e2e)
    for s in range(min(args.warmup, 1)):
        step_e2e(s)
    barrier()
    cc.reset_coder_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    h2d = d2h = 0
    step_wall_e2e = []
    for s in range(args.warmup, total):
        tw = time.perf_counter()
        a, b = step_e2e(s)
        step_wall_e2e.append(round((time.perf_counter() - tw) * 1e3, 2))
        h2d += a
        d2h += b
    e1.record()
    barrier()
    t_e2e = max(e0.elapsed_time(e1) * 1e-3, time.perf_counter() - t0)
    h2d += cc.coder_stats["h2d_bytes"]
    d2h += cc.coder_stats["d2h_bytes"]

    # ---- max over ranks
    # one frame at a time (no pipelining): the codec's latency, for reference
    lat = []
    for _ in range(3):
        torch.cuda.synchronize()
        tw = time.perf_counter()
        s_, sh_, k_, c_ = model.compress(*dev_inputs[0])
        out_ = model.decompress(coordinates=c_, strings=s_, shape=sh_, k=k_)
        torch.cuda.synchronize()
        lat.append(round((time.perf_counter() - tw) * 1e3, 2))
    gpu_result = (s_, sh_, k_, c_.cpu().numpy(), out_.cpu().numpy())
    # the interval the reference itself times (utils.py:448-464): compress(path=...) + decompress(path=...) through
    # the latent-coordinate side channel and the container, one frame at a time, file on tmpfs
    file_variant = None
    if rank == 0:
        fdir = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
        fpath = os.path.join(fdir, f"pcc_b200_bench_{os.getpid()}.bin")
        try:
            ft = []
            for _ in range(3):
                torch.cuda.synchronize()
                tw = time.perf_counter()
                model.compress(*dev_inputs[0], path=fpath)
                tm = time.perf_counter()
                fout = model.decompress(path=fpath)
                torch.cuda.synchronize()
                ft.append((round((tm - tw) * 1e3, 2), round((time.perf_counter() - tm) * 1e3, 2)))
            fbytes = os.path.getsize(fpath)
            from pcc_b200.bitstream import gpcc_encode
            side = len(gpcc_encode(c_))
            file_variant = {"compress_ms": [a for a, _ in ft], "decompress_ms": [b for _, b in ft],
                            "points_per_s": dev_inputs[0][0].shape[0] / (min(a + b for a, b in ft) * 1e-3),
                            "file_bytes": fbytes, "bpp": fbytes * 8.0 / dev_inputs[0][0].shape[0],
                            "side_channel_bytes": side, "side_channel_bytes_per_latent_voxel": side / c_.shape[0],
                            "decoded_equals_in_memory": bool(torch.equal(fout, out_)),
                            "note": "one frame at a time (no pipelining), ColorModel.compress(x, Q, path) + "
                                    ".decompress(path): own octree side channel instead of the reference's tmc3 "
                                    "subprocess, container of model/model.py:214-315; file on tmpfs"}
        finally:
            if os.path.exists(fpath):
                os.unlink(fpath)
    gc.enable()
    pipe.close()
    (t_dev, t_e2e), npts_all = pcc_b200.frames.reduce_throughput([t_dev, t_e2e], npts, dev)
    # configs[4] as a secondary record of every multi-GPU run: ONE 12-bit ~5 M-point scene cut into x-slabs over the
    # same ranks (halo exchange per conv layer over NCCL, strong scaling), verified against one GPU
    scene_line = None
    if world > 1 and not args.no_scene_pass:
        try:
            scene_line = measure_scene(args, rank, world, local_rank, model, steps=2, warmup=1, verify=True)
        except Exception as e:  # the headline line must survive a failure of the secondary measurement
            scene_line = {"error": f"{type(e).__name__}: {e}"[:300]}
    if rank != 0:
        return

    roof = None
    if not args.no_roofline:
        roof, _ = roofline_pass(model, *dev_inputs[0], peaks, clocks, args.dump_layers)
        if roof is not None and world == 1:
            # every non-dense kernel class of the path (hashing, kernel map, stride / expand / sort, top-k,
            # quantise + index, gathers, small-channel convolutions) on this frame's maps: algorithmic bytes
            # (SURVEY.md 8(d) per-unit figures) / CUDA-event time, as a fraction of the measured HBM copy rate
            try:
                from tools.class_bench import measure_classes
                roof["classes"] = [
                    {k: (round(v, 4) if isinstance(v, float) else v) for k, v in c.items() if k != "kernels"}
                    for c in measure_classes(device=str(dev), radius=args.radius, grid=args.grid, reps=3,
                                             hbm_gbs=peaks.get("hbm_gbs"))]
                roof["classes_peak"] = {"hbm_gbs": peaks.get("hbm_gbs", 6650.0),
                                        "source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"}
            except Exception as e:
                roof["classes_error"] = f"{type(e).__name__}: {e}"[:200]
    tf32 = None
    if world == 1 and args.mode == "exact" and not args.no_tf32_pass:
        tf32 = tf32_pass(args, model, dev, dev_inputs[0], B, peaks, clocks)
    base = verified = None
    if world == 1 and not args.no_cpu_baseline:
        base, verified = cpu_baseline(args, args.config, gpu_result)

    n_frame = get(0).shape[0]
    line = {
        "metric": METRIC, "value": npts_all / t_dev, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32" if args.mode == "exact" else "tf32",
        "data": "synthetic",
        "config": {"workload": workload_name(args, n_frame),
                   "mode": args.mode, "frames_per_step_per_gpu": B, "frames_in_flight_per_gpu": min(args.depth, B),
                   "parallelism": f"frame-parallel x{world}, {min(args.depth, B)}-deep frame pipeline per GPU, scheduler "
                                  f"'{args.scheduler}' (CPU range coder and count read-backs of one frame overlap the "
                                  "kernels of the others)",
                   "l2": "per-step working set (>1 GB of activations and kernel maps) exceeds the 126 MB L2; no flush",
                   "includes_cpu_range_coder": True},
        "e2e": {"value": npts_all / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d // args.steps,
                "d2h_bytes_per_step": d2h // args.steps, "ms_per_step": t_e2e / args.steps * 1e3},
        "gpu_launches": launches,
        "clocks": clocks,
        "coder": {"encode_ms_per_frame": coder["encode_s"] / (args.steps * B) * 1e3,
                  "decode_ms_per_frame": coder["decode_s"] / (args.steps * B) * 1e3,
                  "copies_ms_per_frame": (coder["d2h_s"] + coder["h2d_s"]) / (args.steps * B) * 1e3,
                  "symbols_per_frame": coder["symbols"] // (args.steps * B),
                  "note": "reference's sequential CPU rANS (rank 0's frames), inside the timed region; with "
                          "several frames in flight it overlaps other frames' kernels"},
        "single_frame_latency_ms": lat,
        "step_host_ms": {"device_resident": step_wall, "e2e": step_wall_e2e, "pipeline": pipe_stats,
                         "note": "host wall clock per step on rank 0 (launch side; the decoded cloud's "
                                 "D2H ends each e2e step)"},
    }
    if roof is not None:
        line["roofline"] = roof
    if tf32 is not None:
        line["tf32"] = tf32
    if base is not None:
        line["cpu_baseline"] = base
    if verified is not None:
        line["verified_bit_exact_vs_oracle"] = verified
    if file_variant is not None:
        line["file_variant"] = file_variant
    if scene_line is not None:
        keep = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "e2e", "gpu_launches",
                "verified_bit_exact_vs_one_gpu", "string_sha1", "error")
        line["scene"] = {k: scene_line[k] for k in keep if k in scene_line}
        if "config" in scene_line:
            line["scene"]["config"] = {k: scene_line["config"][k] for k in ("workload", "parallelism", "points_in_out_per_rank")}
    print(json.dumps(line))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product path has no CPU fallback); "
                         "use --impl reference for the CPU restatement")
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    try:
        (run_scene if args.scene else run_ours)(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
