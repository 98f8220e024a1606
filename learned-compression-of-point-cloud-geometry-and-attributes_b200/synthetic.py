"""Deterministic synthetic inputs: voxelised spherical shells with smooth + noisy
RGB, and the uniform quality map of utils.compress_model_ours (utils.py:436-445).  Datasets and
checkpoints are unavailable offline."""
import numpy as np
import torch


def shell_points(radius, center=None, grid=1024):
    """Integer voxels p with  radius - 0.5 <= |p - c| < radius + 0.5, lexicographic order."""
    c = np.array([grid // 2] * 3 if center is None else center, dtype=np.int64)
    R = int(np.ceil(radius + 1))
    ax = np.arange(-R, R + 1, dtype=np.int64)
    dx, dy = np.meshgrid(ax, ax, indexing="ij")
    d2 = (dx * dx + dy * dy).reshape(-1)
    lo, hi = (radius - 0.5) ** 2, (radius + 0.5) ** 2
    # |z| in [zmin, zmax]:  lo <= d2 + z^2 < hi
    zmax = np.ceil(np.sqrt(np.maximum(hi - d2, 0.0))).astype(np.int64) - 1
    zmin = np.ceil(np.sqrt(np.maximum(lo - d2, 0.0))).astype(np.int64)
    valid = (hi - d2 > 0) & (zmax >= zmin)
    dxv, dyv = dx.reshape(-1)[valid], dy.reshape(-1)[valid]
    zmin, zmax = zmin[valid], zmax[valid]
    cnt = zmax - zmin + 1
    idx = np.repeat(np.arange(cnt.shape[0]), cnt)
    off = np.arange(cnt.sum()) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    zpos = zmin[idx] + off
    px, py = dxv[idx], dyv[idx]
    pts_pos = np.stack([px, py, zpos], axis=1)
    neg = zpos > 0
    pts_neg = np.stack([px[neg], py[neg], -zpos[neg]], axis=1)
    pts = np.concatenate([pts_pos, pts_neg], axis=0) + c[None, :]
    order = np.lexsort((pts[:, 2], pts[:, 1], pts[:, 0]))
    return pts[order].astype(np.int32)


def shell_frame(radius=252.0, center=None, grid=1024, seed=0):
    """-> x [N, 6] float32 torch tensor (CPU): xyz voxel ints as float, rgb in [0, 1]."""
    pts = shell_points(radius, center, grid)
    rng = np.random.default_rng(seed)
    u = pts.astype(np.float64) / float(grid)
    rgb = np.stack([0.5 + 0.4 * np.sin(6.0 * u[:, 0] + 2.0 * u[:, 1]),
                    0.5 + 0.4 * np.cos(5.0 * u[:, 1] - 3.0 * u[:, 2]),
                    0.5 + 0.4 * np.sin(4.0 * u[:, 2] + 1.0 * u[:, 0])], axis=1)
    rgb = np.clip(rgb + rng.uniform(-0.05, 0.05, size=rgb.shape), 0.0, 1.0)
    return torch.from_numpy(np.concatenate([pts.astype(np.float32), rgb.astype(np.float32)], axis=1))


def uniform_q_map(x, q_g=0.5, q_a=0.5):
    """(coordinates [N,4] float, features [N,2]) — channel 0 = geometry, 1 = attribute
    (utils.py:438-440)."""
    N = x.shape[0]
    coords = torch.cat([torch.zeros((N, 1), dtype=x.dtype), x[:, :3]], dim=1)
    feats = torch.cat([torch.full((N, 1), float(q_g)), torch.full((N, 1), float(q_a))], dim=1)
    return coords, feats


def sequence_frame(t, n_frames=300, seed=0):
    """Frame t of the synthetic sequence of configs[3]: radius 252 + 4 sin(2 pi t / n), drifting
    centre, colours seeded with t."""
    r = 252.0 + 4.0 * np.sin(2.0 * np.pi * t / n_frames)
    c = [512 + int(round(8 * np.sin(2.0 * np.pi * t / n_frames))), 512, 512 + (t % 7) - 3]
    return shell_frame(r, c, 1024, seed=seed + t)
