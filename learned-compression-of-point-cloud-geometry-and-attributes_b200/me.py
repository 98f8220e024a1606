"""MinkowskiEngine-shaped module surface over the B200 kernels.

Exports exactly the names and signatures /root/reference/model/{model,transforms,blocks,
entropy_models}.py and utils.py use from ``import MinkowskiEngine as ME``:
``SparseTensor`` (.C, .F, .device, .tensor_stride, .coordinate_manager,
.features_at_coordinates), ``MinkowskiConvolution``, ``MinkowskiGenerativeConvolutionTranspose``,
``MinkowskiConvolutionTranspose``, ``MinkowskiPruning``, ``MinkowskiReLU``, ``MinkowskiLeakyReLU``.
Modules are ``nn.Module``s with parameters named ``kernel`` ([K, Cin, Cout]; [Cin, Cout] for
kernel_size 1) and ``bias`` ([1, Cout]) so state_dict keys and shapes match ME's.

Semantics follow SURVEY.md Appendix A.  Row order: a SparseTensor built from user coordinates
keeps the given row order (as ME does); maps created by strided / transposed convolutions come
out in canonical lexicographic (b,x,y,z) order (ME's order there is unspecified, §7.3 H2).
"""
import math

import torch
import torch.nn as nn

from . import ops
from ._lib import PccError


class CoordinateManager:
    """Identity object shared by tensors that may reuse each other's maps; remembers the map
    created for each tensor stride so a non-generative transposed convolution can find it."""

    def __init__(self):
        self.maps = {}

    def register(self, cmap):
        self.maps.setdefault(cmap.ts, cmap)


def _as_stride(ts):
    if isinstance(ts, (list, tuple)):
        return int(ts[0])
    if torch.is_tensor(ts):
        return int(ts.reshape(-1)[0].item())
    return int(ts)


class SparseTensor:
    """ME.SparseTensor (72 call sites on the path)."""

    def __init__(self, features=None, coordinates=None, tensor_stride=1, device=None,
                 coordinate_manager=None, coordinate_map_key=None, *, _cmap=None, _manager=None):
        if _cmap is not None:  # internal fast path: map already built, no validation
            self._F = features
            self.cmap = _cmap
            self.coordinate_manager = _manager if _manager is not None else CoordinateManager()
            self.coordinate_manager.register(_cmap)
            return
        if features is None or coordinates is None:
            raise PccError("SparseTensor needs features= and coordinates=")
        if device is None:
            device = coordinates.device if coordinates.is_cuda else features.device
        device = torch.device(device)
        if device.type != "cuda":
            raise PccError("SparseTensor must live on a CUDA device: there is no CPU fallback")
        coords = coordinates.to(device)
        if coords.is_floating_point():
            coords = torch.floor(coords)  # ME floors non-integer coordinates (model/model.py:184-188)
        coords = coords.to(torch.int32).contiguous()
        feats = features.to(device=device, dtype=torch.float32).contiguous()
        if coords.dim() != 2 or coords.shape[1] != 4 or feats.shape[0] != coords.shape[0]:
            raise PccError(f"bad shapes: coordinates {tuple(coords.shape)}, features {tuple(feats.shape)}")
        ts = _as_stride(tensor_stride)
        keys, status = ops.pack_coords(coords)
        skeys, _ = ops.sort_keys(keys, want_perm=False)
        dup = (skeys[1:] == skeys[:-1]).any() if keys.shape[0] > 1 else torch.zeros((), dtype=torch.bool, device=device)
        bad, has_dup = torch.stack([status[0] != 0, dup.reshape(())]).tolist()
        if bad:
            raise PccError("coordinate out of the packable range (|x|,|y|,|z| < 32768, 0 <= b < 65536)")
        if has_dup:
            raise PccError("duplicate coordinates are not supported (the codec never passes any)")
        self._F = feats
        self.cmap = ops.CoordMap(keys, ts, False)
        self.cmap._coords = coords
        self.coordinate_manager = coordinate_manager if coordinate_manager is not None else CoordinateManager()
        self.coordinate_manager.register(self.cmap)

    # -- attributes the reference reads
    @property
    def F(self):
        return self._F

    @property
    def C(self):
        return self.cmap.coords

    @property
    def device(self):
        return self._F.device

    @property
    def tensor_stride(self):
        return [self.cmap.ts] * 3

    @property
    def shape(self):
        return self._F.shape

    def __len__(self):
        return self._F.shape[0]

    def features_at_coordinates(self, query_coordinates):
        """On-grid queries: exact gather-or-zero."""
        q = query_coordinates.to(self.device)
        if q.is_floating_point():
            q = torch.floor(q)
        qkeys, _ = ops.pack_coords(q.to(torch.int32))
        rows = self.cmap.lookup(qkeys)
        from . import autograd as AG
        return AG.gather(self._F, rows) if AG.wants_grad(self._F) else ops.gather_rows(self._F, rows)


def _replace(x, feats, cmap=None):
    return SparseTensor(feats, _cmap=cmap if cmap is not None else x.cmap, _manager=x.coordinate_manager)


class _ConvBase(nn.Module):
    is_transpose = False
    generative = False

    def __init__(self, in_channels, out_channels, kernel_size=-1, stride=1, dilation=1, bias=False,
                 kernel_generator=None, expand_coordinates=False, convolution_mode=None, dimension=None):
        super().__init__()
        if dimension not in (None, 3):
            raise PccError("only dimension=3 is supported")
        if dilation != 1 or kernel_generator is not None:
            raise PccError("dilation / custom kernel generators are not on the codec path")
        if kernel_size not in (1, 2, 3) or stride not in (1, 2):
            raise PccError("kernel_size must be 1, 2 or 3 and stride 1 or 2")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride = kernel_size, stride
        self.kernel_volume = kernel_size ** 3
        if self.kernel_volume > 1:
            self.kernel = nn.Parameter(torch.empty(self.kernel_volume, in_channels, out_channels))
        else:
            self.kernel = nn.Parameter(torch.empty(in_channels, out_channels))
        self.bias = nn.Parameter(torch.empty(1, out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        """ME's default initialisation."""
        n = (self.out_channels if self.is_transpose else self.in_channels) * self.kernel_volume
        stdv = 1.0 / math.sqrt(n)
        with torch.no_grad():
            self.kernel.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)

    def _maps(self, x):
        """-> (out_map, kernel_map)"""
        cm = x.cmap
        if not self.is_transpose:
            if self.stride == 1:
                out_map = cm
            else:
                out_map = cm.stride2()
            return out_map, cm.kernel_map(out_map, self.kernel_size, transpose=False)
        if self.stride != 2:
            raise PccError("transposed convolution is implemented for stride 2")
        existing = None if self.generative else x.coordinate_manager.maps.get(cm.ts // 2)
        out_map = existing if existing is not None else cm.expand(self.kernel_size)
        return out_map, cm.kernel_map(out_map, self.kernel_size, transpose=True)

    def forward(self, input, coordinates=None, *, act=ops.ACT_NONE):
        if coordinates is not None:
            raise PccError("explicit output coordinates are not on the codec path")
        out_map, kmap = self._maps(input)
        from . import autograd as AG
        if self.training and AG.wants_grad(input.F, self.kernel):
            # train() mode under autograd (train.py:189-191): ME's ConvolutionBackward is the data gradient
            # over the reversed kernel map + pcc_conv_wgrad (autograd.py)
            feats = AG.sparse_conv(input.F, input.cmap, out_map, self.kernel_size, self.kernel, self.bias,
                                   transpose=self.is_transpose, act=act)
        else:
            feats = ops.conv(input.F, kmap, self.kernel, self.bias, act=act)
        return _replace(input, feats, out_map)

    def extra_repr(self):
        return (f"in={self.in_channels}, out={self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}, bias={self.bias is not None}")


class MinkowskiConvolution(_ConvBase):
    """ME.MinkowskiConvolution (57 sites)."""


class MinkowskiConvolutionTranspose(_ConvBase):
    """ME.MinkowskiConvolutionTranspose (model/entropy_models.py:298,302): reuses the manager's
    map at the finer stride when one exists, else generates it."""
    is_transpose = True


class MinkowskiGenerativeConvolutionTranspose(_ConvBase):
    """ME.MinkowskiGenerativeConvolutionTranspose (model/blocks.py:84; model/entropy_models.py:286,290)."""
    is_transpose = True
    generative = True


class MinkowskiReLU(nn.Module):
    def __init__(self, inplace=False):
        super().__init__()

    def forward(self, input):
        return _replace(input, torch.relu(input.F))


class MinkowskiLeakyReLU(nn.Module):
    def __init__(self, negative_slope=0.01, inplace=False):
        super().__init__()
        self.negative_slope = negative_slope

    def forward(self, input):
        return _replace(input, torch.nn.functional.leaky_relu(input.F, self.negative_slope))


class MinkowskiPruning(nn.Module):
    """ME.MinkowskiPruning (model/blocks.py:90,126): keep rows whose mask is True, order kept."""

    def forward(self, input, mask):
        rows = torch.nonzero(mask.to(input.device).reshape(-1)).reshape(-1).to(torch.int32)
        keys = ops.take_keys(input.cmap.keys, rows)
        from . import autograd as AG
        feats = AG.gather(input.F, rows) if AG.wants_grad(input.F) else ops.gather_rows(input.F, rows)
        cmap = ops.CoordMap(keys, input.cmap.ts, input.cmap.is_sorted)
        return SparseTensor(feats, _cmap=cmap, _manager=CoordinateManager())
