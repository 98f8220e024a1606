"""Quality maps: the user-facing control of the codec.

The codec is conditioned per point by Q = [q_g, q_a] in [0, 1] (geometry, attribute).  The reference
builds Q next to the call (utils.py:436-445) and, for the view-dependent experiments, from per-point
scores (evaluate_view_dep.py:209-216 gradient along an axis, :252-258 region of interest behind a
plane).  These helpers build the same (coordinates [N,4], features [N,2]) pair, on whatever device
the points live on; `ColorModel.compress(x, Q)` takes it as a tuple.
"""
import torch


def _coords(points):
    """[N,3] or [N,>=3] (xyz first) -> [N,4] with a zero batch column (utils.py:438)."""
    xyz = points[:, :3]
    return torch.cat([torch.zeros((xyz.shape[0], 1), dtype=xyz.dtype, device=xyz.device), xyz], dim=1)


def q_map(points, q_g, q_a):
    """utils.py:436-445.  q_g, q_a: floats (uniform map) or per-point arrays/tensors [N] / [N,1]."""
    n, dev = points.shape[0], points.device

    def col(q):
        if isinstance(q, (int, float)):
            return torch.ones((n, 1), device=dev) * float(q)
        t = torch.as_tensor(q, device=dev).reshape(-1, 1).to(torch.float32)
        if t.shape[0] != n:
            raise ValueError(f"quality map has {t.shape[0]} entries for {n} points")
        return t

    return _coords(points), torch.cat([col(q_g), col(q_a)], dim=1).to(torch.float32)


def view_scores(points, axis, lo, hi):
    """evaluate_view_dep.py:212: clip((p[axis] - min) / (max - min), 0, 1), in float64 like numpy."""
    return torch.clamp((points[:, axis].to(torch.float64) - lo) / (hi - lo), 0.0, 1.0)


def roi_scores(points, axis, plane):
    """evaluate_view_dep.py:254: 0 where p[axis] < plane, else 1."""
    return torch.where(points[:, axis] < plane, 0.0, 1.0).to(torch.float64)


def view_q_map(points, q_g, q_a, axis, lo, hi):
    """View-dependent map (evaluate_view_dep.py:209-216): quality falls off linearly along `axis`."""
    s = view_scores(points, axis, lo, hi)
    return q_map(points, q_g * s, q_a * s)


def roi_q_map(points, q_g, q_a, axis, plane):
    """Region-of-interest map (evaluate_view_dep.py:252-258): full quality at and beyond `plane`."""
    s = roi_scores(points, axis, plane)
    return q_map(points, q_g * s, q_a * s)


Q_MAP_CONFIG = {"lambda_A_min": 0, "lambda_A_max": 6400, "lambda_G_min": 0, "lambda_G_max": 200,
                "mode": "exponential"}  # configs/Ours.yaml:38-43


def lambda_map(q_feats, config=None):
    """Training-time loss weights of a quality map (data/q_map.py:188-205, ``scale_q_map``):
    column 0 (geometry) and 1 (attribute) of Q in [0, 1] -> lambda in [lambda_min, lambda_max]."""
    import math
    cfg = Q_MAP_CONFIG if config is None else config
    out = q_feats.clone().to(torch.float32)
    if cfg["mode"] == "exponential":
        a_A, b_A = math.log2(cfg["lambda_A_max"] + cfg["lambda_A_min"]), cfg["lambda_A_min"] - 1
        a_G, b_G = math.log2(cfg["lambda_G_max"] + cfg["lambda_G_min"]), cfg["lambda_G_min"] - 1
        out[:, 0] = 2 ** (out[:, 0] * a_G) + b_G
        out[:, 1] = 2 ** (out[:, 1] * a_A) + b_A
    elif cfg["mode"] == "quadratic":
        a_A, b_A = cfg["lambda_A_max"] - cfg["lambda_A_min"], cfg["lambda_A_min"]
        a_G, b_G = cfg["lambda_G_max"] - cfg["lambda_G_min"], cfg["lambda_G_min"]
        out[:, 0] = out[:, 0] ** 2 * a_G + b_G
        out[:, 1] = out[:, 1] ** 2 * a_A + b_A
    else:
        raise ValueError("Unknown Q_map mode")
    return out
