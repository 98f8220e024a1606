"""Training losses of the reference (loss.py) over the B200 operators.

Same class names, config keys and call signature ``loss(gt, pred) -> (total, {id: value})`` as
/root/reference/loss.py:7-67.  ``gt`` is a SparseTensor-like (``.C`` [N,4] int with batch ids,
``.F`` [N,3] colours); ``pred`` is the dictionary ``ColorModel.forward`` returns.  Coordinate
membership (``torch.isin`` over flattened coordinates, loss.py:121-127,166-171) is a hash probe of
the packed keys; ``features_at_coordinates`` is the on-grid gather; ``MinkowskiAvgPooling``
(loss.py:150-151) is the sparse convolution with per-offset identity weights divided by the
neighbour count.  ``ColorSSIM`` (loss.py:197-452) is not built: no shipped config selects it.
"""
import math

import torch

from . import autograd as AG
from . import ops
from ._lib import PccError


def _as_map(t):
    """SparseTensor-like / internal ST / CoordMap -> (CoordMap in canonical order, features or None)."""
    if isinstance(t, ops.CoordMap):
        return t, None
    m = getattr(t, "m", None)
    if isinstance(m, ops.CoordMap):
        return m, t.F
    cm = getattr(t, "cmap", None)
    if isinstance(cm, ops.CoordMap) and cm.is_sorted:
        return cm, t.F
    ts = getattr(t, "tensor_stride", 1)
    ts = ts[0] if isinstance(ts, (list, tuple)) else ts
    C = t.C
    if not C.is_cuda:
        raise PccError("loss inputs must live on the CUDA device (no CPU fallback)")
    m, perm = ops.canonical_map(C.int(), int(ts))
    F = None if t.F is None else AG.gather(t.F.float().to(C.device), perm.to(torch.int32))
    return m, F


def _member(query_map, ref_map):
    """bool [n_query]: coordinate of query present in ref (torch.isin on flattened coordinates)."""
    return ref_map.lookup(query_map.keys) >= 0


def _at(src_map, src_F, query_map):
    """src.features_at_coordinates(query.C) for on-grid queries."""
    return AG.gather(src_F, src_map.lookup(query_map.keys)) if src_map is not query_map else src_F


def avg_pool(F, in_map, out_map, kernel_size=3):
    """ME.MinkowskiAvgPooling(kernel_size, stride): mean of the inputs present in the window of each
    output coordinate (0 where the window is empty)."""
    c = F.shape[1]
    K = kernel_size ** 3
    w = torch.eye(c + 1, device=F.device).unsqueeze(0).repeat(K, 1, 1).contiguous()
    x = torch.cat([F, torch.ones((F.shape[0], 1), device=F.device)], dim=1)
    kmap = in_map.kernel_map(out_map, kernel_size, transpose=False)
    s = ops.conv(x.contiguous(), kmap, w, None, mode=ops.MODE_EXACT, round_out=False)
    n = s[:, c:]
    return torch.where(n > 0, s[:, :c] / n.clamp(min=1.0), torch.zeros_like(s[:, :c]))


class Loss:
    """loss.py:7-67."""

    def __init__(self, config):
        self.losses = {}
        table = {"BPPLoss": BPPLoss, "ColorLoss": ColorLoss, "FocalLoss": FocalLoss,
                 "Multiscale_FocalLoss": Multiscale_FocalLoss}
        for id, setting in config.items():
            setting = dict(setting)
            setting["id"] = id
            key = setting["type"]
            if key in table:
                self.losses[id] = table[key](setting)
            elif key == "ColorSSIM":
                raise PccError("ColorSSIM is not built (no shipped config selects it)")
            else:
                print("Not found {}".format(key))

    def __call__(self, gt, pred):
        total_loss = 0
        losses = {}
        gt_map = _as_map(gt)
        for _, loss in self.losses.items():
            item = loss(gt, pred, _gt=gt_map)
            losses[loss.identifier] = item
            total_loss = total_loss + item
        return total_loss, losses


class BPPLoss:
    """loss.py:69-88: sum(log likelihood) / (-ln 2 * num_points) * weight."""

    def __init__(self, config):
        self.weight, self.identifier, self.key = config["weight"], config["id"], config["key"]

    def __call__(self, gt, pred, _gt=None):
        likelihoods = pred["likelihoods"][self.key]
        num_points = gt.C.shape[0]
        loss = 0.0
        for likelihood in likelihoods:
            loss = loss + torch.log(likelihood).sum() / (-math.log(2) * num_points)
        return loss.mean() * self.weight


class ColorLoss:
    """loss.py:91-112: L1/L2 at the ground-truth voxels, weighted by channel 1 of the q-map."""

    def __init__(self, config):
        self.identifier = config["id"]
        self.l1 = config["loss"] == "L1"

    def __call__(self, gt, pred, _gt=None):
        gt_map, gt_F = _gt if _gt is not None else _as_map(gt)
        p_map, p_F = _as_map(pred["prediction"])
        q_map, q_F = _as_map(pred["q_map"])
        pred_colors = _at(p_map, p_F, gt_map)
        d = gt_F - pred_colors
        loss = d.abs() if self.l1 else d * d
        loss = loss * _at(q_map, q_F, gt_map)[:, 1].unsqueeze(1)
        return loss.mean()


def _focal(logit, inside, alpha, gamma):
    p_z = torch.sigmoid(logit)
    pt_z = torch.where(inside, p_z, 1 - p_z)
    alpha_z = torch.where(inside, torch.full_like(p_z, alpha), torch.full_like(p_z, 1 - alpha))
    pt_z = torch.clip(pt_z, 1e-2, 1)
    return -alpha_z * (1 - pt_z) ** gamma * torch.log(pt_z)


class FocalLoss:
    """loss.py:115-141 (single-scale variant; reads pred["lambdas"])."""

    def __init__(self, config):
        self.identifier, self.alpha, self.gamma = config["id"], config["alpha"], config["gamma"]

    def __call__(self, gt, pred, _gt=None):
        gt_map, _ = _gt if _gt is not None else _as_map(gt)
        p_map, p_F = _as_map(pred["prediction"])
        focal = _focal(p_F[:, 0] + 0.5, _member(p_map, gt_map), self.alpha, self.gamma)
        return focal.mean() * pred["lambdas"][0][0]


class Multiscale_FocalLoss:
    """loss.py:143-195: focal loss of every up-sampling stage's occupancy logits against the
    ground-truth occupancy of that scale, weighted by the 3x3x3-average of q-map channel 0."""

    def __init__(self, config):
        self.identifier, self.alpha, self.gamma = config["id"], config["alpha"], config["gamma"]

    def __call__(self, gt, pred, _gt=None):
        predictions = list(reversed(pred["occ_predictions"]))  # finest scale first (:158-159)
        points = list(reversed(pred["points"]))
        q_map, q_F = _as_map(pred["q_map"])
        q_F = q_F.detach()
        loss = 0.0
        for prediction, coords in zip(predictions, points):
            p_map, p_F = _as_map(prediction)
            c_map, _ = _as_map(coords)
            focal = _focal(p_F[:, 0], _member(p_map, c_map), self.alpha, self.gamma)
            q_avgs = avg_pool(q_F, q_map, p_map)            # self.pooling(q_map, coordinates=prediction.C)
            nxt = q_map.stride2()
            q_F, q_map = avg_pool(q_F, q_map, nxt), nxt      # q_map = self.down_pooling(q_map)
            loss = loss + (focal * q_avgs[:, 0]).mean()
        return loss
