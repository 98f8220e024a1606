"""ORACLE — TEST INFRASTRUCTURE ONLY.

Stand-in ``MinkowskiEngine`` / ``compressai`` / ``open3d`` / ``bitstream`` modules that let the
reference's OWN, UNMODIFIED ``model/model.py``, ``model/transforms.py``, ``model/blocks.py``,
``model/entropy_models.py`` and ``utils.py`` (imported from /root/reference at fixture-generation
time) execute on the CPU over the oracle primitives (``oracle.prims`` / ``oracle.entropy``).

What this pins: the reference's control flow — layer order, channel widths, which tensor feeds
which layer, the FiLM / residual / top-k / pruning / sorting logic, the state_dict names of every
module in model/*.py — is executed as written, so golden vectors made this way
(tests/golden/make_golden.py) are outputs of the reference code itself.
What it does NOT pin: the semantics of the third-party packages (SURVEY.md Appendix A/B,
[upstream-recalled]); those are supplied by the oracle primitives both here and in oracle.codec.

Row order follows ME: a SparseTensor keeps the row order it is given; a stride-1 convolution keeps
its input's order; strided / transposed convolutions create canonical (lexicographic) maps.
"""
import sys
import types

import numpy as np
import torch
import torch.nn as nn

from . import prims
from .entropy import EntropyBottleneckOracle, GaussianConditionalOracle, get_scale_table


# ------------------------------------------------------------------------------ MinkowskiEngine
class CoordinateManager:
    pass


class SparseTensor:
    """ME.SparseTensor."""

    def __init__(self, features=None, coordinates=None, tensor_stride=1, device=None,
                 coordinate_manager=None, coordinate_map_key=None):
        c = coordinates.detach().cpu()
        if c.is_floating_point():
            c = torch.floor(c)
        self._C = c.to(torch.int32).contiguous()
        self._F = features.detach().cpu().to(torch.float32).contiguous()
        assert self._C.shape[0] == self._F.shape[0]
        ts = tensor_stride
        if isinstance(ts, (list, tuple)):
            ts = ts[0]
        self._ts = int(ts)
        self.coordinate_manager = coordinate_manager if coordinate_manager is not None else CoordinateManager()
        keys = prims.pack_keys(self._C.numpy())
        assert np.unique(keys).size == keys.size, "duplicate coordinates"

    @property
    def C(self):
        return self._C

    @property
    def F(self):
        return self._F

    @property
    def tensor_stride(self):
        return [self._ts] * 3

    @property
    def device(self):
        return self._F.device

    @property
    def shape(self):
        return self._F.shape

    def _ot(self):
        """-> (canonical oracle tensor, perm) with canonical = given[perm]."""
        c = self._C.numpy()
        perm = np.argsort(prims.pack_keys(c), kind="stable")
        return prims.OT(c[perm], self._F.numpy()[perm], self._ts, _sorted=True), perm

    def features_at_coordinates(self, q):
        ot, _ = self._ot()
        return torch.from_numpy(prims.features_at(ot, q.detach().cpu().numpy()))


def _from_ot(ot, manager=None):
    return SparseTensor(features=torch.from_numpy(ot.F.copy()), coordinates=torch.from_numpy(ot.C.copy()),
                        tensor_stride=ot.ts, coordinate_manager=manager)


class _ConvBase(nn.Module):
    transposed = False
    generative = False

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, convolution_mode=None, dimension=None):
        super().__init__()
        assert dimension == 3 and dilation == 1
        self.in_channels, self.out_channels = int(in_channels), int(out_channels)
        self.kernel_size, self.stride = int(kernel_size), int(stride)
        K = self.kernel_size ** 3
        shape = (K, self.in_channels, self.out_channels) if K > 1 else (self.in_channels, self.out_channels)
        self.kernel = nn.Parameter(torch.empty(shape).uniform_(-0.1, 0.1))
        self.bias = nn.Parameter(torch.empty(1, self.out_channels).uniform_(-0.1, 0.1)) if bias else None

    def forward(self, input):
        ot, perm = input._ot()
        W = self.kernel.detach().cpu().numpy()
        b = None if self.bias is None else self.bias.detach().cpu().numpy()
        if self.transposed:
            # Appendix A5: without a finer map in the same coordinate manager (always the case on
            # the reference path) the non-generative transpose generates its map like the
            # generative one does.
            out = prims.conv_transpose(ot, W, b, self.kernel_size, self.stride)
            return _from_ot(out)
        out = prims.conv(ot, W, b, self.kernel_size, self.stride)
        if self.stride == 1:  # same coordinate map: rows stay in the input's order
            F = np.empty_like(out.F)
            F[perm] = out.F
            return SparseTensor(features=torch.from_numpy(F), coordinates=input.C, tensor_stride=input._ts,
                                coordinate_manager=input.coordinate_manager)
        return _from_ot(out)


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
        return SparseTensor(features=torch.from_numpy(prims.relu(x.F.numpy())), coordinates=x.C,
                            tensor_stride=x._ts, coordinate_manager=x.coordinate_manager)


class MinkowskiLeakyReLU(nn.Module):
    def __init__(self, negative_slope=0.01, inplace=False):
        super().__init__()
        self.negative_slope = negative_slope

    def forward(self, x):
        return SparseTensor(features=torch.from_numpy(prims.leaky_relu(x.F.numpy(), self.negative_slope)),
                            coordinates=x.C, tensor_stride=x._ts, coordinate_manager=x.coordinate_manager)


# ------------------------------------------------------------------------------ compressai
class _LowerBound(nn.Module):
    def __init__(self, bound):
        super().__init__()
        self.register_buffer("bound", torch.Tensor([float(bound)]))


class _EntropyModel(nn.Module):
    def __init__(self, likelihood_bound=1e-9):
        super().__init__()
        self.likelihood_lower_bound = _LowerBound(likelihood_bound)
        self.register_buffer("_offset", torch.IntTensor())
        self.register_buffer("_quantized_cdf", torch.IntTensor())
        self.register_buffer("_cdf_length", torch.IntTensor())

    def _publish(self, o):
        self._offset = torch.from_numpy(o.offset.copy())
        self._quantized_cdf = torch.from_numpy(np.ascontiguousarray(o.quantized_cdf, np.int32))
        self._cdf_length = torch.from_numpy(o.cdf_length.copy())


class EntropyBottleneck(_EntropyModel):
    """compressai.entropy_models.EntropyBottleneck over EntropyBottleneckOracle (Appendix B1)."""

    def __init__(self, channels, *args, tail_mass=1e-9, init_scale=10, filters=(3, 3, 3, 3), **kwargs):
        super().__init__()
        f = (1,) + tuple(filters) + (1,)
        scale = init_scale ** (1 / (len(filters) + 1))
        for i in range(len(filters) + 1):
            m = torch.full((channels, f[i + 1], f[i]), float(np.log(np.expm1(1 / scale / f[i + 1]))))
            self.register_parameter(f"_matrix{i}", nn.Parameter(m))
            self.register_parameter(f"_bias{i}", nn.Parameter(torch.empty(channels, f[i + 1], 1).uniform_(-0.5, 0.5)))
            if i < len(filters):
                self.register_parameter(f"_factor{i}", nn.Parameter(torch.zeros(channels, f[i + 1], 1)))
        self.quantiles = nn.Parameter(torch.tensor([-float(init_scale), 0.0, float(init_scale)]).repeat(channels, 1, 1))
        target = np.log(2 / tail_mass - 1)
        self.register_buffer("target", torch.Tensor([-target, 0, target]))
        self._o = None

    def _oracle(self):
        if self._o is None:
            names = [n for n, _ in self.named_parameters()]
            self._o = EntropyBottleneckOracle({n: getattr(self, n).detach().cpu().numpy() for n in names})
        return self._o

    def update(self, force=False):
        self._o = None
        self._publish(self._oracle())
        return True

    def compress(self, x):
        return [self._oracle().compress(x[0].detach().cpu().numpy())]

    def decompress(self, strings, size):
        return torch.from_numpy(self._oracle().decompress(strings[0], int(size[0]))).unsqueeze(0)

    def forward(self, x, training=None):
        z_hat, lik = self._oracle().forward(x[0].detach().cpu().numpy())
        return torch.from_numpy(z_hat).unsqueeze(0), torch.from_numpy(lik).unsqueeze(0)

    def loss(self):
        return torch.zeros(())


class GaussianConditional(_EntropyModel):
    """compressai.entropy_models.GaussianConditional over GaussianConditionalOracle (Appendix B2)."""

    def __init__(self, scale_table, *args, scale_bound=0.11, tail_mass=1e-9, **kwargs):
        super().__init__()
        self.lower_bound_scale = _LowerBound(scale_bound)
        self.register_buffer("scale_table", torch.Tensor())
        self.register_buffer("scale_bound", torch.Tensor([float(scale_bound)]))
        self._o = None

    def update_scale_table(self, scale_table, force=False):
        self._o = GaussianConditionalOracle()
        assert np.array_equal(np.asarray(scale_table, np.float32), self._o.scale_table)
        self.scale_table = torch.from_numpy(self._o.scale_table.copy())
        self._publish(self._o)
        return True

    def build_indexes(self, scales):
        return torch.from_numpy(self._o.build_indexes(scales.detach().cpu().numpy()))

    def compress(self, inputs, indexes, means=None):
        return [self._o.compress(inputs[0].detach().cpu().numpy(), indexes[0].numpy(), means[0].detach().cpu().numpy())]

    def decompress(self, strings, indexes, dtype=torch.float, means=None):
        out = self._o.decompress(strings[0], indexes[0].numpy(), means[0].detach().cpu().numpy())
        return torch.from_numpy(out).unsqueeze(0)

    def forward(self, inputs, scales, means=None, training=None):
        y_hat, lik = self._o.forward(inputs[0].detach().cpu().numpy(), scales[0].detach().cpu().numpy(),
                                     means[0].detach().cpu().numpy())
        return torch.from_numpy(y_hat).unsqueeze(0), torch.from_numpy(lik).unsqueeze(0)


class CompressionModel(nn.Module):
    """compressai.models.base.CompressionModel.update / aux_loss (model/model.py:15,34,45)."""

    def update(self, scale_table=None, force=False):
        if scale_table is None:
            scale_table = get_scale_table()
        for m in self.modules():
            if isinstance(m, EntropyBottleneck):
                m.update(force=force)
            if isinstance(m, GaussianConditional):
                m.update_scale_table(scale_table, force=force)
        return True

    def aux_loss(self):
        return sum(m.loss() for m in self.modules() if isinstance(m, EntropyBottleneck))


class _NonNegativeParametrizer(nn.Module):
    def __init__(self, minimum=0.0, reparam_offset=2 ** -18):
        super().__init__()
        pedestal = reparam_offset ** 2
        self.register_buffer("pedestal", torch.Tensor([pedestal]))
        self.lower_bound = nn.Module()
        self.lower_bound.register_buffer("bound", torch.Tensor([(minimum + pedestal) ** 0.5]))


class GDN(nn.Module):
    """compressai.layers.GDN: parameters only (model/blocks.py:27 constructs it, nothing calls it)."""

    def __init__(self, in_channels, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        self.inverse = bool(inverse)
        self.beta_reparam = _NonNegativeParametrizer(minimum=beta_min)
        self.beta = nn.Parameter(torch.sqrt(torch.ones(in_channels) + self.beta_reparam.pedestal))
        self.gamma_reparam = _NonNegativeParametrizer()
        self.gamma = nn.Parameter(torch.sqrt(gamma_init * torch.eye(in_channels) + self.gamma_reparam.pedestal))

    def forward(self, x):
        raise RuntimeError("GDN is never called on the reference path")


# ------------------------------------------------------------------------------ installation
def install():
    """Put the stand-in modules into sys.modules (only names that are not importable for real)."""
    me = types.ModuleType("MinkowskiEngine")
    for k in ("SparseTensor", "CoordinateManager", "MinkowskiConvolution", "MinkowskiConvolutionTranspose",
              "MinkowskiGenerativeConvolutionTranspose", "MinkowskiPruning", "MinkowskiReLU", "MinkowskiLeakyReLU"):
        setattr(me, k, globals()[k])
    ca = types.ModuleType("compressai")
    ca_em = types.ModuleType("compressai.entropy_models")
    ca_em.EntropyBottleneck, ca_em.GaussianConditional = EntropyBottleneck, GaussianConditional
    ca_models = types.ModuleType("compressai.models")
    ca_base = types.ModuleType("compressai.models.base")
    ca_base.CompressionModel = CompressionModel
    ca_layers = types.ModuleType("compressai.layers")
    ca_layers.GDN = GDN
    ca.entropy_models, ca.models, ca.layers = ca_em, ca_models, ca_layers
    ca_models.base = ca_base
    o3d = types.ModuleType("open3d")       # imported by model/model.py:8 and utils.py:2, unused on the path
    bs = types.ModuleType("bitstream")     # model/model.py:9, used only by the file container
    bs.BitStream = type("BitStream", (), {})
    for name, mod in [("MinkowskiEngine", me), ("compressai", ca), ("compressai.entropy_models", ca_em),
                      ("compressai.models", ca_models), ("compressai.models.base", ca_base),
                      ("compressai.layers", ca_layers), ("open3d", o3d), ("bitstream", bs)]:
        sys.modules.setdefault(name, mod)


def load_reference_model(config, reference_root="/root/reference"):
    """Import the reference's model package as it lies on disk and build ColorModel(config)."""
    install()
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    import importlib
    model_pkg = importlib.import_module("model.model")
    return model_pkg.ColorModel(config)
