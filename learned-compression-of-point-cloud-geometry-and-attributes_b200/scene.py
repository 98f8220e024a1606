"""One large scene coded by several GPUs (BASELINE.json configs[4]).

`compress_scene` / `decompress_scene` are ColorModel.compress / .decompress (model/model.py:95-208)
for a scene cut into x-slabs (slab.py): every rank passes the points of its own slab, the analysis
and synthesis transforms run on the slab with a one-voxel halo exchanged per layer, the stride-8
latent is gathered for the (single, sequential) entropy coder, and top-k pruning selects scene-wide.
The returned strings / shape / k / coordinates / decoded cloud are bit-identical to what one GPU
produces for the whole scene (tests/test_gpu_slab.py).
"""
import torch

from . import ops
from ._lib import PccError
from .codec import ST, stage
from .slab import Slab, SlabMap, slab_bounds, DistComm, ThreadComm, ALIGN  # noqa: F401


def split_scene(x, Q, bounds):
    """Host helper: cut (x [N,6], Q = (coords [N,4], feats [N,2])) into len(bounds) - 1 slabs by x."""
    qc, qf = Q
    out = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        mx = (x[:, 0] >= a) & (x[:, 0] < b)
        mq = (qc[:, 1] >= a) & (qc[:, 1] < b)
        out.append((x[mx], (qc[mq], qf[mq])))
    return out


def balanced_bounds(x, world):
    """Slab boundaries (multiples of 32) balancing the point counts of x [N,6] over `world` slabs."""
    xs = x[:, 0].long()
    lo = (int(xs.min()) // ALIGN) * ALIGN
    cells = torch.bincount((xs - lo) // ALIGN)
    return slab_bounds(int(xs.min()), int(xs.max()), world, cells.tolist())


def _restrict(st, slab):
    """The rows of a whole-scene ST that fall into the slab, as an ST over a SlabMap."""
    a, b = slab.row_range(st.m.keys, slab.x0, slab.x1)
    m = SlabMap(st.m.keys[a:b].clone(), st.m.ts, slab)
    return ST(m.lift(st.F[a:b]), m)


@torch.no_grad()
def compress_scene(model, x_own, Q_own, slab):
    """x_own [n,6], Q_own: this rank's slab of the scene.  -> (strings, shape, k, coordinates) of the
    WHOLE scene (identical on every rank)."""
    if model.entropy_model_map is not None:
        raise PccError("the slab partition supports the MeanScaleHyperprior_Map configuration")
    comm = slab.comm
    with stage("scene.enc.inputs"):
        inp, Qt = model._inputs(x_own, Q_own)
        m = SlabMap(inp.m.keys, 1, slab)
        Qm = m if Qt.m is inp.m else SlabMap(Qt.m.keys, 1, slab)
        inp, Qt = ST(m.lift(inp.F), m), ST(Qm.lift(Qt.F), Qm)
    with stage("scene.enc.g_a"):
        y, Qy, k = model.g_a(inp, Qt)
    with stage("scene.enc.gather"):
        keys_all, _ = comm.all_gather_rows(y.m.keys)
        F_all, _ = comm.all_gather_rows(y.F[:y.m.n].contiguous())
        y_full = ST(F_all, ops.CoordMap(keys_all, y.m.ts, True))
    points, strings, shape = model.entropy_model.compress(y_full)
    return strings, shape, k, y_full.m.coords


@torch.no_grad()
def decompress_scene(model, coordinates, strings, shape, k, slab, gather=True):
    """-> decoded cloud [N,6] of the whole scene (gather=True, identical on every rank) or of this
    rank's slab."""
    dev = model.device
    coordinates = coordinates.to(dev)
    if coordinates.shape[1] == 3:
        coordinates = torch.cat([torch.zeros((coordinates.shape[0], 1), device=dev, dtype=coordinates.dtype),
                                 coordinates], dim=1)
    if coordinates.is_floating_point():
        coordinates = torch.floor(coordinates)
    with stage("scene.dec.entropy"):
        y_map, _ = ops.canonical_map(coordinates.int(), 8)
        z_map = y_map.stride2().stride2()
        # the latent is small (stride 8): every rank decodes it, then keeps its slab
        y_hat, Q_hat = model.entropy_model.decompress([y_map, z_map], strings, shape)
        y_s, Q_s = _restrict(y_hat, slab), _restrict(Q_hat, slab)
    with stage("scene.dec.g_s"):
        x_hat = model.g_s(y_s, Q_s, k=k)
    with stage("scene.dec.finalize"):
        out = ops.color_finalize(x_hat.m.keys, x_hat.F[:x_hat.m.n].contiguous())
        if gather:
            out, _ = slab.comm.all_gather_rows(out)
    return out
