#!/usr/bin/env python
"""Training-step timing: forward + loss + backward + Adam on one synthetic batch
shaped like the reference's (configs/Ours.yaml: batch_size 8 of 128^3 blocks, train.py:169-203).

  python tools/train_bench.py [--batch 8] [--radius 56] [--steps 5] [--warmup 3] [--layers out.csv]

Prints one JSON line: points/s through the step, ms per phase (CUDA events on the launching
stream), and the weight-gradient kernel's share and useful TFLOP/s (2 * pairs * cin * cout per
launch over the event-timed launch durations, layers with cin, cout >= 64) against the FP32 FMA peak.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

LOSS_CONFIG = {
    "Multiscale_FocalLoss": {"type": "Multiscale_FocalLoss", "alpha": 0.5, "gamma": 2.0},
    "ColorLoss": {"type": "ColorLoss", "loss": "L2"},
    "bpp-y": {"type": "BPPLoss", "key": "y", "weight": 1.0},
    "bpp-z": {"type": "BPPLoss", "key": "z", "weight": 1.0},
}


class T:
    def __init__(self, C, F):
        self.C, self.F = C, F


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--radius", type=float, default=56.0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--layers", default=None)
    ap.add_argument("--mode", default="exact", choices=["exact", "tf32"],
                    help="tf32: forward and data gradient of the >= 32-channel layers on tcgen05 (weight gradient fp32)")
    ap.add_argument("--host-profile", default=None, help="write a cProfile table (with callers) of one step here")
    ap.add_argument("--torch-profile", default=None, help="write a torch.profiler table of one step here")
    a = ap.parse_args()
    import pcc_b200 as pcc
    dev = "cuda:0"
    pcc.ops.set_default_mode(pcc.ops.MODE_TF32 if a.mode == "tf32" else pcc.ops.MODE_EXACT)
    model = pcc.ColorModel(pcc.OURS_CONFIG)
    pcc.init_synthetic(model, 0)
    model.to(dev).train()
    cs, fs, qs = [], [], []
    for b in range(a.batch):
        x = pcc.synthetic.shell_frame(a.radius - b, center=[64] * 3, grid=128, seed=b)
        n = x.shape[0]
        cs.append(torch.cat([torch.full((n, 1), b, dtype=torch.int32), x[:, :3].int()], 1))
        fs.append(x[:, 3:6].float())
        qs.append(torch.stack([torch.full((n,), (b % 4) / 3.0), torch.full((n,), ((b + 1) % 3) / 2.0)], 1))
    coords, colours, q = torch.cat(cs).to(dev), torch.cat(fs).to(dev), torch.cat(qs).to(dev)
    lam = pcc.qmap.lambda_map(q)
    x, Q, Lam = T(coords, colours), T(coords, q), T(coords, lam)
    loss_fn = pcc.loss.Loss(LOSS_CONFIG)
    main_p = [p for n, p in model.named_parameters() if not n.endswith(".quantiles")]
    aux_p = [p for n, p in model.named_parameters() if n.endswith(".quantiles")]
    opt = torch.optim.Adam(main_p, lr=1e-4)
    aux_opt = torch.optim.Adam(aux_p, lr=1e-3)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def step(marks=None):
        opt.zero_grad(set_to_none=True)
        aux_opt.zero_grad(set_to_none=True)
        out = model(x, Q, Lam)
        total, parts = loss_fn(x, out)
        if marks is not None:
            marks[1].record()
        total.backward()
        if marks is not None:
            marks[2].record()
        torch.nn.utils.clip_grad_norm_(main_p, 1.0)
        opt.step()
        aux = model.aux_loss()
        aux.backward()
        aux_opt.step()
        if os.environ.get("PCC_TRAIN_VERBOSE"):
            print({k: round(float(v.detach()), 4) for k, v in parts.items()}, "aux", round(float(aux.detach()), 2), file=sys.stderr)
        return float(total.detach())

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    l0 = pcc.ops.launches()
    fwd = bwd = opt_ms = 0.0
    t0, t3 = ev(), ev()
    t0.record()
    per = []
    for _ in range(a.steps):
        m = [ev(), ev(), ev(), ev()]
        m[0].record()
        loss = step(m)
        m[3].record()
        per.append(m)
    t3.record()
    torch.cuda.synchronize()
    launches = pcc.ops.launches() - l0
    total_ms = t0.elapsed_time(t3) / a.steps
    fwd = sum(m[0].elapsed_time(m[1]) for m in per) / a.steps
    bwd = sum(m[1].elapsed_time(m[2]) for m in per) / a.steps
    opt_ms = sum(m[2].elapsed_time(m[3]) for m in per) / a.steps
    step_ms = [round(m[0].elapsed_time(m[3]), 1) for m in per]
    if a.host_profile:
        import cProfile, pstats, io
        pr = cProfile.Profile()
        pr.enable()
        step()
        torch.cuda.synchronize()
        pr.disable()
        buf = io.StringIO()
        st = pstats.Stats(pr, stream=buf).sort_stats("tottime")
        st.print_stats(40)
        st.print_callers("tolist|item|nonzero|synchronize|_local_scalar|masked_select|index|unique")
        with open(a.host_profile, "w") as f:
            f.write(buf.getvalue())
    if a.torch_profile:
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        with open(a.torch_profile, "w") as f:
            f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
            f.write("\n\n")
            f.write(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=30, max_name_column_width=70))
    # one profiled step: per-launch CUDA events around every convolution / weight-gradient launch
    pcc.ops.start_conv_profile()
    step()
    recs = pcc.ops.stop_conv_profile()
    big = [r for r in recs if r["mode"] in (-1, -2) and r["cin"] >= 64 and r["cout"] >= 64]
    tc = [r for r in recs if r["mode"] == 1 and r["cin"] >= 64 and r["cout"] >= 64]
    wg_all = sum(r["ms"] for r in recs if r["mode"] < 0)
    conv_all = sum(r["ms"] for r in recs if r["mode"] >= 0)
    wg_tc = any(r["mode"] == -2 for r in big)
    flops = sum(2.0 * r["pairs"] * r["cin"] * r["cout"] for r in big)
    ms = sum(r["ms"] for r in big)
    peak = 148 * 128 * 2 * 1.965e9 / 1e12
    try:
        tc_peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                              "MEASURED_PEAKS.json")))["bf16_tflops"] / 2.0
        tc_src = "tf32 dense = measured bf16 burst / 2 (MEASURED_PEAKS.json)"
    except (OSError, KeyError, ValueError):
        tc_peak, tc_src = 1590.0 / 2.0, "fallback 1590/2"
    if a.layers:
        with open(a.layers, "w") as f:
            f.write("kind,cin,cout,K,n_in,n_out,pairs,ms,tflops\n")
            for r in recs:
                tf = 2.0 * r["pairs"] * r["cin"] * r["cout"] / max(r["ms"], 1e-9) / 1e9
                f.write(f"{'wgrad' if r['mode'] == -1 else 'wgrad_tf32' if r['mode'] == -2 else 'conv'},{r['cin']},{r['cout']},{r['K']},{r['n_in']},"
                        f"{r['n_out']},{r['pairs']},{r['ms']:.4f},{tf:.3f}\n")
    n = int(coords.shape[0])
    print(json.dumps({
        "metric": "training step points/sec (forward + loss + backward + Adam)", "value": n / (total_ms / 1e3),
        "unit": "points/s", "ms_per_step": total_ms, "points_per_batch": n, "batch": a.batch,
        "step_ms": step_ms, "reserved_gb": round(torch.cuda.memory_reserved() / 2**30, 2),
        "phases_ms": {"forward+loss": fwd, "backward": bwd, "clip+adam+aux": opt_ms},
        "gpu_launches_per_step": launches / a.steps, "last_loss": loss, "dtype": "f32" if a.mode == "exact" else "tf32",
        "tensor_core_conv": None if not tc else {
            "kernel": "conv_tf32_kernel (forward + data gradient)", "launches": len(tc),
            "achieved": sum(2.0 * r["pairs"] * r["cin"] * r["cout"] for r in tc) / sum(r["ms"] for r in tc) / 1e9,
            "unit": "TFLOP/s"},
        "conv_ms_per_step": {"forward+data-gradient (pcc_conv_fwd)": conv_all, "weight-gradient (pcc_conv_wgrad)": wg_all},
        "roofline": ({"bound": "fp32_fma", "kernel": "conv_wgrad_kernel", "achieved": flops / max(ms, 1e-9) / 1e9,
                      "peak": peak, "unit": "TFLOP/s", "frac": flops / max(ms, 1e-9) / 1e9 / peak,
                      "launches": len(big), "peak_source": "148 SM x 128 FMA x 2 x 1965 MHz"} if not wg_tc else
                     {"bound": "tensor", "kernel": "wgrad_tf32_kernel", "achieved": flops / max(ms, 1e-9) / 1e9,
                      "peak": tc_peak, "unit": "TFLOP/s", "frac": flops / max(ms, 1e-9) / 1e9 / tc_peak,
                      "launches": len(big), "peak_source": tc_src}),
        "config": {"mode": a.mode, "workload": f"Ours.yaml random-init, batch of {a.batch} synthetic 7-bit shell blocks (radius "
                               f"{a.radius}..{a.radius - a.batch + 1} in 128^3), loss section of configs/Ours.yaml"}}))


if __name__ == "__main__":
    main()
