"""ORACLE — TEST INFRASTRUCTURE ONLY.  **parity unpinned** at the MinkowskiEngine / CompressAI boundary.

Differentiable (torch CPU autograd) stand-ins of the ``MinkowskiEngine`` / ``compressai`` names the
reference's TRAINING path touches, so that the reference's OWN, UNMODIFIED ``ColorModel.forward``
(model/model.py:51-93), ``MeanScaleHyperprior_Map.forward`` (model/entropy_models.py:309-337) and
``loss.py`` run here and ``loss.backward()`` yields gradients for every parameter
(tests/golden/make_golden_train.py).  Coordinates and kernel maps come from the oracle primitives
(``oracle.prims``); features stay torch tensors on the tape:

  * convolution       out = sum_k index_add(F[nbr_k] @ W[k])            (Appendix A2-A5)
  * pruning           F[mask]                                             (Appendix A6)
  * features_at       gather-or-zero at on-grid queries                   (Appendix A7)
  * avg pooling       window mean over the inputs present                 (loss.py:150-151)
  * EntropyBottleneck / GaussianConditional forward in training mode: additive U(-.5,.5) noise,
    CompressAI 1.2.4's likelihood expressions and LowerBound's gradient rule (Appendix B1-B2,
    [upstream-recalled]).

The noise is drawn by ``noise_like`` from a seeded CPU generator so the product can be fed the
same draw.  Only tests / fixture scripts import this module.
"""
import sys
import types

import numpy as np
import torch
import torch.nn as nn

from . import prims
from . import ref_shims as RS

_gen = torch.Generator().manual_seed(0)
noise_log = []  # tensors drawn since the last reseed, in call order (the product replays them)


def reseed(seed):
    _gen.manual_seed(int(seed))
    noise_log.clear()


def noise_like(t):
    n = torch.empty(t.shape, dtype=torch.float32).uniform_(-0.5, 0.5, generator=_gen)
    noise_log.append(n)
    return n.to(t.dtype)


class CoordinateManager:
    pass


class SparseTensor:
    def __init__(self, features=None, coordinates=None, tensor_stride=1, device=None,
                 coordinate_manager=None, coordinate_map_key=None):
        c = coordinates.detach().cpu()
        if c.is_floating_point():
            c = torch.floor(c)
        self._C = c.to(torch.int32).contiguous()
        self._F = features.to(torch.get_default_dtype())  # fixture scripts may run the tape in float64
        assert self._C.shape[0] == self._F.shape[0]
        ts = tensor_stride[0] if isinstance(tensor_stride, (list, tuple)) else tensor_stride
        self._ts = int(ts)
        self.coordinate_manager = coordinate_manager if coordinate_manager is not None else CoordinateManager()
        self._keys = prims.pack_keys(self._C.numpy())
        assert np.unique(self._keys).size == self._keys.size, "duplicate coordinates"

    C = property(lambda self: self._C)
    F = property(lambda self: self._F)
    tensor_stride = property(lambda self: [self._ts] * 3)
    device = property(lambda self: self._F.device)
    shape = property(lambda self: self._F.shape)

    def _canonical(self):
        perm = np.argsort(self._keys, kind="stable")
        return self._keys[perm], perm

    def features_at_coordinates(self, q):
        keys, perm = self._canonical()
        qn = q.detach().cpu().numpy()
        assert (qn == np.floor(qn)).all()
        rows = prims.lookup(np.ascontiguousarray(keys), qn.astype(np.int32))
        hit = torch.from_numpy(rows >= 0)
        src = torch.from_numpy(perm[np.maximum(rows, 0)].astype(np.int64))
        return self._F[src] * hit.unsqueeze(1).to(self._F.dtype)


def _conv_apply(F_can, nbr, W, bias, n_out):
    if W.dim() == 2:
        W = W.unsqueeze(0)
    W = W.to(F_can.dtype)
    out = torch.zeros((n_out, W.shape[2]), dtype=F_can.dtype)
    for k in range(nbr.shape[0]):
        o = np.nonzero(nbr[k] >= 0)[0]
        if o.size:
            out = out.index_add(0, torch.from_numpy(o.astype(np.int64)),
                                F_can[torch.from_numpy(nbr[k][o].astype(np.int64))] @ W[k])
    return out if bias is None else out + bias.to(out.dtype)


class _ConvBase(RS._ConvBase):
    def forward(self, input):
        keys, perm = input._canonical()
        keys = np.ascontiguousarray(keys)
        C_can = input._C.numpy()[perm]
        F_can = input._F[torch.from_numpy(perm.astype(np.int64))]
        ts = input._ts
        if self.transposed:
            ts_out = ts // 2
            out_coords = prims.expand_coords(C_can, self.kernel_size, ts_out)
            offs = -prims.kernel_offsets(self.kernel_size) * ts_out
        elif self.stride == 1:
            ts_out, out_coords = ts, C_can
            offs = prims.kernel_offsets(self.kernel_size) * ts
        else:
            ts_out = ts * 2
            out_coords = prims.stride_coords(C_can, ts_out)
            offs = prims.kernel_offsets(self.kernel_size) * ts
        nbr = prims.kernel_map(keys, out_coords, offs)
        out = _conv_apply(F_can, nbr, self.kernel, self.bias, out_coords.shape[0])
        if not self.transposed and self.stride == 1:  # same map: rows stay in the input's order
            inv = np.empty_like(perm)
            inv[perm] = np.arange(perm.size)
            return SparseTensor(features=out[torch.from_numpy(inv.astype(np.int64))], coordinates=input.C,
                                tensor_stride=ts, coordinate_manager=input.coordinate_manager)
        return SparseTensor(features=out, coordinates=torch.from_numpy(out_coords.copy()), tensor_stride=ts_out)


class MinkowskiConvolution(_ConvBase):
    pass


class MinkowskiConvolutionTranspose(_ConvBase):
    transposed = True


class MinkowskiGenerativeConvolutionTranspose(_ConvBase):
    transposed = True
    generative = True


class MinkowskiPruning(nn.Module):
    def forward(self, x, mask):
        mask = mask.detach().cpu().bool()
        return SparseTensor(features=x.F[mask], coordinates=x.C[mask], tensor_stride=x._ts)


class MinkowskiReLU(nn.Module):
    def __init__(self, inplace=False):
        super().__init__()

    def forward(self, x):
        return SparseTensor(features=torch.relu(x.F), coordinates=x.C, tensor_stride=x._ts,
                            coordinate_manager=x.coordinate_manager)


class MinkowskiLeakyReLU(nn.Module):
    def __init__(self, negative_slope=0.01, inplace=False):
        super().__init__()
        self.negative_slope = negative_slope

    def forward(self, x):
        return SparseTensor(features=torch.nn.functional.leaky_relu(x.F, self.negative_slope), coordinates=x.C,
                            tensor_stride=x._ts, coordinate_manager=x.coordinate_manager)


class MinkowskiAvgPooling(nn.Module):
    """Window mean over the inputs present; ``coordinates=`` fixes the output coordinates."""

    def __init__(self, kernel_size=-1, stride=1, dilation=1, kernel_generator=None, dimension=None):
        super().__init__()
        self.kernel_size, self.stride = int(kernel_size), int(stride)

    def forward(self, x, coordinates=None):
        keys, perm = x._canonical()
        F_can = x._F[torch.from_numpy(perm.astype(np.int64))]
        ts = x._ts
        if coordinates is not None:
            assert self.stride == 1
            out_coords, ts_out = coordinates.detach().cpu().numpy().astype(np.int32), ts
        elif self.stride == 1:
            out_coords, ts_out = x._C.numpy()[perm], ts
        else:
            ts_out = ts * self.stride
            out_coords = prims.stride_coords(x._C.numpy()[perm], ts_out)
        nbr = prims.kernel_map(np.ascontiguousarray(keys), out_coords, prims.kernel_offsets(self.kernel_size) * ts)
        c = F_can.shape[1]
        eye = torch.eye(c + 1, dtype=F_can.dtype).unsqueeze(0).repeat(nbr.shape[0], 1, 1)
        s = _conv_apply(torch.cat([F_can, torch.ones(F_can.shape[0], 1, dtype=F_can.dtype)], dim=1), nbr, eye, None,
                        out_coords.shape[0])
        n = s[:, c:]
        out = torch.where(n > 0, s[:, :c] / n.clamp(min=1.0), torch.zeros_like(s[:, :c]))
        return SparseTensor(features=out, coordinates=torch.from_numpy(np.ascontiguousarray(out_coords)),
                            tensor_stride=ts_out)


# ------------------------------------------------------------------------------ compressai (training)
class _LowerBoundFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bound):
        ctx.save_for_backward(x, bound)
        return torch.max(x, bound)

    @staticmethod
    def backward(ctx, g):
        x, bound = ctx.saved_tensors
        return ((x >= bound) | (g < 0)).to(g.dtype) * g, None


class _LowerBound(nn.Module):
    def __init__(self, bound):
        super().__init__()
        self.register_buffer("bound", torch.Tensor([float(bound)]))

    def forward(self, x):
        return _LowerBoundFn.apply(x, self.bound)


class EntropyBottleneck(RS.EntropyBottleneck):
    def __init__(self, channels, *args, **kwargs):
        super().__init__(channels, *args, **kwargs)
        self.channels = channels
        self.likelihood_lower_bound = _LowerBound(1e-9)

    def _logits_cumulative(self, v, stop_gradient=False):
        p = (lambda t: t.detach()) if stop_gradient else (lambda t: t)
        logits = v
        for i in range(5):
            logits = torch.matmul(torch.nn.functional.softplus(p(getattr(self, f"_matrix{i}"))), logits)
            logits = logits + p(getattr(self, f"_bias{i}"))
            if i < 4:
                logits = logits + torch.tanh(p(getattr(self, f"_factor{i}"))) * torch.tanh(logits)
        return logits

    def forward(self, x, training=None):
        training = self.training if training is None else training
        if not training:
            return super().forward(x)
        values = x.permute(1, 0, 2).reshape(self.channels, 1, -1)
        outputs = values + noise_like(values)
        lower = self._logits_cumulative(outputs - 0.5)
        upper = self._logits_cumulative(outputs + 0.5)
        sign = -torch.sign(lower + upper).detach()
        lik = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))
        lik = self.likelihood_lower_bound(lik)
        shape = (self.channels, x.shape[0], -1)
        return outputs.reshape(shape).permute(1, 0, 2), lik.reshape(shape).permute(1, 0, 2)

    def loss(self):
        # compressai 1.2.4: _logits_cumulative(self.quantiles, stop_gradient=True) — only the quantiles learn
        return torch.abs(self._logits_cumulative(self.quantiles, stop_gradient=True) - self.target).sum()


class GaussianConditional(RS.GaussianConditional):
    def __init__(self, scale_table, *args, scale_bound=0.11, **kwargs):
        super().__init__(scale_table, *args, scale_bound=scale_bound, **kwargs)
        self.lower_bound_scale = _LowerBound(scale_bound)
        self.likelihood_lower_bound = _LowerBound(1e-9)

    def forward(self, inputs, scales, means=None, training=None):
        training = self.training if training is None else training
        if not training:
            return super().forward(inputs, scales, means)
        outputs = inputs + noise_like(inputs)
        values = outputs - means if means is not None else outputs
        s = self.lower_bound_scale(scales)
        values = torch.abs(values)
        cum = lambda t: 0.5 * torch.erfc(-(2 ** -0.5) * t)
        lik = cum((0.5 - values) / s) - cum((-0.5 - values) / s)
        return outputs, self.likelihood_lower_bound(lik)


class CompressionModel(RS.CompressionModel):
    def update(self, scale_table=None, force=False):
        if scale_table is None:
            scale_table = RS.get_scale_table()
        for m in self.modules():
            if isinstance(m, RS.EntropyBottleneck):
                m.update(force=force)
            if isinstance(m, RS.GaussianConditional):
                m.update_scale_table(scale_table, force=force)
        return True

    def aux_loss(self):
        return sum(m.loss() for m in self.modules() if isinstance(m, EntropyBottleneck))


NAMES = ("MinkowskiEngine", "compressai", "compressai.entropy_models", "compressai.models",
         "compressai.models.base", "compressai.layers", "open3d", "bitstream")


def install():
    """Replace whatever stand-ins are installed with the differentiable ones."""
    me = types.ModuleType("MinkowskiEngine")
    for k in ("SparseTensor", "CoordinateManager", "MinkowskiConvolution", "MinkowskiConvolutionTranspose",
              "MinkowskiGenerativeConvolutionTranspose", "MinkowskiPruning", "MinkowskiReLU", "MinkowskiLeakyReLU",
              "MinkowskiAvgPooling"):
        setattr(me, k, globals()[k])
    ca = types.ModuleType("compressai")
    ca_em = types.ModuleType("compressai.entropy_models")
    ca_em.EntropyBottleneck, ca_em.GaussianConditional = EntropyBottleneck, GaussianConditional
    ca_models = types.ModuleType("compressai.models")
    ca_base = types.ModuleType("compressai.models.base")
    ca_base.CompressionModel = CompressionModel
    ca_layers = types.ModuleType("compressai.layers")
    ca_layers.GDN = RS.GDN
    ca.entropy_models, ca.models, ca.layers = ca_em, ca_models, ca_layers
    ca_models.base = ca_base
    o3d = types.ModuleType("open3d")
    bs = types.ModuleType("bitstream")
    bs.BitStream = type("BitStream", (), {})
    for name, mod in [("MinkowskiEngine", me), ("compressai", ca), ("compressai.entropy_models", ca_em),
                      ("compressai.models", ca_models), ("compressai.models.base", ca_base),
                      ("compressai.layers", ca_layers), ("open3d", o3d), ("bitstream", bs)]:
        sys.modules[name] = mod


def load_reference(config, reference_root="/root/reference"):
    """-> (ColorModel(config), loss module) of the reference, imported as they lie on disk."""
    install()
    for m in [m for m in sys.modules if m == "model" or m.startswith("model.") or m in ("utils", "loss")]:
        del sys.modules[m]
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    import importlib
    model_pkg = importlib.import_module("model.model")
    loss_mod = importlib.import_module("loss")
    return model_pkg.ColorModel(config), loss_mod
