"""The codec's encode/decode path (ColorModel.compress / .decompress) on the B200 kernels.

Host-side mirror of /root/reference/model/{model,transforms,blocks,entropy_models}.py: the same
class names, constructor configs, module tree and therefore state_dict keys / parameter shapes,
so a checkpoint of the reference loads unchanged.  The forward passes are not layer-by-layer
translations: each block runs as a fused plan over ``ops`` —
  * one coordinate map per resolution level, shared by every layer (and by the x and Q branches)
    instead of a fresh coordinate manager per ``ME.SparseTensor(...)`` call;
  * ReLU / LeakyReLU, the FiLM modulation ``F * beta + gamma`` (model/blocks.py:37-40) and the
    residual add (model/blocks.py:49-52) live in the convolution epilogue;
  * result-preserving removals of SURVEY.md §7.3 H4: occupancy head evaluated for channel 0 only,
    q_up evaluated only at the surviving coordinates, the last h_s layer evaluated only at y's
    coordinates, z_hat taken from the quantiser instead of decoding the string just encoded.
Each method cites the reference lines it follows.
"""
import os

import torch
import torch.nn as nn

from . import autograd as AG
from . import ops
from . import me as ME
from .compressai_compat import CompressionModel, EntropyBottleneck, GaussianConditional, Wait, drive
from ._lib import PccError
from .slab import SlabMap

RELU, LEAKY, NONE = ops.ACT_RELU, ops.ACT_LEAKY, ops.ACT_NONE

# Optional stage instrumentation, off on the hot path:
#  * STAGE_MS (tools/stage_profile.py): when set to a dict, every `with stage(name)` block is bracketed by device
#    synchronisations and its wall time accumulated;
#  * PCC_NVTX=1 in the environment: every stage is an NVTX range (torch.cuda.nvtx), so an `ncu --nvtx` / Nsight
#    Systems capture shows g_a, h_a, the coders, h_s and the three up-sampling stages by name.
STAGE_MS = None
NVTX = os.environ.get("PCC_NVTX", "0") == "1"


class stage:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if NVTX:
            torch.cuda.nvtx.range_push(self.name)
        if STAGE_MS is not None:
            import time
            torch.cuda.synchronize()
            self.t0 = time.perf_counter()

    def __exit__(self, *exc):
        if STAGE_MS is not None:
            import time
            torch.cuda.synchronize()
            STAGE_MS[self.name] = STAGE_MS.get(self.name, 0.0) + (time.perf_counter() - self.t0) * 1e3
        if NVTX:
            torch.cuda.nvtx.range_pop()
        return False


def _conv(cin, cout, k=3, stride=1, bias=True):
    return ME.MinkowskiConvolution(in_channels=cin, out_channels=cout, kernel_size=k, stride=stride,
                                   bias=bias, dimension=3)


class ST:
    """Internal sparse tensor of the fused path: features + the level's shared CoordMap."""
    __slots__ = ("F", "m")

    def __init__(self, F, m):
        self.F, self.m = F, m

    @property
    def C(self):
        return self.m.coords

    @property
    def tensor_stride(self):
        return [self.m.ts] * 3

    @property
    def device(self):
        return self.F.device

    def features_at_coordinates(self, q):
        return self.to_me().features_at_coordinates(q)

    def to_me(self):
        return ME.SparseTensor(self.F, _cmap=self.m)


def apply_conv(x, layer, out_map=None, transpose=False, **epilogue):
    """One fused convolution: x (ST) -> ST on out_map (default: same map)."""
    out_map = x.m if out_map is None else out_map
    if layer.training and AG.wants_grad(x.F, layer.kernel, epilogue.get("film"), epilogue.get("residual")):
        # training (model/model.py:51-93 under autograd): same operator, epilogue unfused
        return ST(AG.sparse_conv(x.F, x.m, out_map, layer.kernel_size, layer.kernel, layer.bias,
                                 transpose=transpose, **epilogue), out_map)
    if isinstance(x.m, SlabMap):  # one slab of a partitioned scene: owned rows computed, halo rows exchanged
        return ST(x.m.conv(x.F, out_map, layer, transpose, epilogue), out_map)
    if x.F.shape[0] != x.m.n:  # the kernels index rows by the map: a shorter matrix would be read out of bounds
        raise PccError(f"feature matrix has {x.F.shape[0]} rows, its coordinate map {x.m.n}")
    kmap = x.m.kernel_map(out_map, layer.kernel_size, transpose=transpose)
    return ST(ops.conv(x.F, kmap, layer.kernel, layer.bias, **epilogue), out_map)


def on_map(m, F):
    """Features of m's rows as a tensor over m: on a SlabMap, owned-row features get their halo rows."""
    if isinstance(m, SlabMap) and F.shape[0] != m.n_local:
        return m.lift(F)
    return F


def gather(F, rows, add=None):
    """out[i] = F[rows[i]] (0 when rows[i] < 0) (+ add[i]); differentiable when a gradient is wanted."""
    if AG.wants_grad(F, add):
        return AG.gather(F, rows, add)
    return ops.gather_rows(F, rows, add=add)


def rows_of(src_map, dst_map):
    """Row of each dst coordinate in src (None when the maps are the same object)."""
    if src_map is dst_map:
        return None
    return src_map.lookup(dst_map.keys)


def seq3(x, seq, acts=(RELU, NONE), **last):
    """Conv, act, Conv[, act]: nn.Sequential(conv, ReLU, conv) with the activations fused."""
    x = apply_conv(x, seq[0], act=acts[0])
    return apply_conv(x, seq[2], act=acts[1], **last)


def seq5(x, seq):
    """Conv, ReLU, Conv, ReLU, Conv (the predict / post / q_pre stacks)."""
    x = apply_conv(x, seq[0], act=RELU)
    x = apply_conv(x, seq[2], act=RELU)
    return apply_conv(x, seq[4])


class _Parametrizer(nn.Module):
    """Carries compressai NonNegativeParametrizer's buffers so GDN state_dict keys line up."""

    def __init__(self, minimum=0.0, reparam_offset=2 ** -18):
        super().__init__()
        pedestal = reparam_offset ** 2
        self.register_buffer("pedestal", torch.Tensor([pedestal]))
        self.lower_bound = nn.Module()
        self.lower_bound.register_buffer("bound", torch.Tensor([(minimum + pedestal) ** 0.5]))


class MinkowskiGDN(nn.Module):
    """model/blocks.py:260-282.  Constructed by ScaledBlock (:27) but never called on the path;
    kept only so parameter names/shapes match."""

    def __init__(self, in_channels, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        self.inverse = bool(inverse)
        self.beta_reparam = _Parametrizer(minimum=beta_min)
        self.beta = nn.Parameter(torch.sqrt(torch.ones(in_channels) + self.beta_reparam.pedestal))
        self.gamma_reparam = _Parametrizer()
        self.gamma = nn.Parameter(torch.sqrt(gamma_init * torch.eye(in_channels) + self.gamma_reparam.pedestal))

    def forward(self, x):
        raise PccError("MinkowskiGDN is never called on the reference path")


class ScaledBlock(nn.Module):
    """model/blocks.py:10-53: conv_2(conv_1(x) * beta + gamma) + x, four k3 convolutions."""

    def __init__(self, N, encode=True, scale=True):
        super().__init__()
        self.encode, self.scale = encode, scale
        self.conv_1 = nn.Sequential(_conv(N, N), ME.MinkowskiReLU(inplace=False), _conv(N, N))
        self.conv_2 = nn.Sequential(_conv(N, N), ME.MinkowskiReLU(inplace=False), _conv(N, N),
                                    ME.MinkowskiReLU(inplace=False))
        self.gdn = MinkowskiGDN(N, inverse=(not self.encode))

    def forward(self, x, condition):
        # FiLM rows: condition.features_at_coordinates(x.C) (:37); residual: x_res at x.C (:50)
        t = seq3(x, self.conv_1, film=condition.F, film_rows=rows_of(condition.m, x.m))
        t = apply_conv(t, self.conv_2[0], act=RELU)
        return apply_conv(t, self.conv_2[2], act=RELU, residual=x.F)


class ConvBlock(nn.Module):
    """model/blocks.py:57-73 (unused by the shipped configs)."""

    def __init__(self, N):
        super().__init__()
        self.conv_1 = nn.Sequential(_conv(N, N), ME.MinkowskiReLU(inplace=False))
        self.conv_2 = nn.Sequential(_conv(N, N), ME.MinkowskiReLU(inplace=False))

    def forward(self, x):
        return apply_conv(apply_conv(x, self.conv_1[0], act=RELU), self.conv_2[0], act=RELU)


class GenerativeUpBlock(nn.Module):
    """model/blocks.py:78-181: generative x2 up-sampling, refinement, occupancy logit, top-k
    pruning with the transmitted point count."""

    def __init__(self, N_in, N_out, predict=False, dense=True, condition_ablation=None):
        super().__init__()
        self.dense, self.condition_ablation, self.predict = dense, condition_ablation, predict
        self.conv = ME.MinkowskiGenerativeConvolutionTranspose(
            in_channels=N_in, out_channels=N_out, kernel_size=3, stride=2, bias=True, dimension=3)
        self.conv_2 = nn.Sequential(_conv(N_out, N_out), ME.MinkowskiReLU(inplace=False), _conv(N_out, N_out))
        self.prune = ME.MinkowskiPruning()
        if self.predict:
            self.occ_predict = nn.Sequential(_conv(N_out, N_out), ME.MinkowskiReLU(inplace=False),
                                             _conv(N_out, N_out))
        self._head = None

    def _occ_head(self):
        """Channel 0 of occ_predict's last convolution: only F[:, 0] is ever read (:142), so the
        other N_out - 1 output channels are not computed (4 kept for 16-byte alignment)."""
        last = self.occ_predict[2]
        ver = (last.kernel._version, last.kernel.data_ptr())
        head = self._head
        if head is None or head[0] != ver:
            with ops.cache_lock:
                c = min(4, last.out_channels)
                w, b = last.kernel.detach()[:, :, :c].contiguous(), last.bias.detach()[:, :c].contiguous()
                ops.publish(w, b)
                head = self._head = (ver, w, b)
        return head[1], head[2]

    def _logits(self, x, full):
        t = apply_conv(x, self.occ_predict[0], act=RELU)
        if full:
            return apply_conv(t, self.occ_predict[2]).F
        last = self.occ_predict[2]
        if last.training and AG.wants_grad(t.F, last.kernel):  # only channel 0 reaches the loss (loss.py:170)
            c = min(4, last.out_channels)
            return AG.sparse_conv(t.F, x.m, x.m, 3, last.kernel[:, :, :c], last.bias[:, :c])
        kmap = x.m.kernel_map(x.m, 3)
        if ops.default_mode() == ops.MODE_TF32 and kmap.n_groups == 1 and not isinstance(x.m, SlabMap):
            # tolerance mode: project onto the 27 head vectors first, then gather scalars (pcc_conv_head0)
            return ops.conv_head0(t.F, kmap, last.kernel.detach(), None if last.bias is None else last.bias.detach())
        w, b = self._occ_head()
        return ops.conv(t.F, kmap, w, b)

    @staticmethod
    def _topk(logits, cand_map, k):
        """_topk_prediction (:130-150): per batch id, the k largest logits of channel 0."""
        if isinstance(cand_map, SlabMap):  # scene-wide k over all slabs
            return cand_map.topk(logits, int(k[0]))
        if cand_map.n_batches is not None:
            nb = cand_map.n_batches if cand_map.n else 0
        else:
            nb = int((cand_map.keys[-1] >> 48).item()) + 1 if cand_map.n else 0
        if len(k) < max(nb, 1):
            raise PccError(f"corrupt stream: {len(k)} point counts for {nb} batch elements")
        if nb <= 1:
            if not 0 <= int(k[0]) <= cand_map.n:
                raise PccError(f"corrupt stream: point count {int(k[0])} exceeds the {cand_map.n} candidate voxels of its stage")
            return ops.topk_rows(logits, int(k[0]))
        bounds = torch.searchsorted(cand_map.keys >> 48, torch.arange(nb + 1, device=logits.device)).tolist()
        for b in range(nb):
            if not 0 <= int(k[b]) <= bounds[b + 1] - bounds[b]:
                raise PccError(f"corrupt stream: point count {int(k[b])} of batch element {b} exceeds its "
                               f"{bounds[b + 1] - bounds[b]} candidate voxels")
        sel = [ops.topk_rows(logits[bounds[b]:bounds[b + 1]], int(k[b])) + bounds[b] for b in range(nb)]
        return torch.cat(sel).to(torch.int32)

    def forward(self, x, coords=None, k=None, full_predictions=False):
        if not self.predict:
            # q_up_i (:179-181): generate then prune to `coords`; evaluated directly at `coords`
            return apply_conv(x, self.conv, out_map=coords, transpose=True)
        cand = x.m.expand(3)
        x = apply_conv(x, self.conv, out_map=cand, transpose=True)
        if self.dense:
            if self.condition_ablation is None:
                x = seq3(x, self.conv_2)
            logits = self._logits(x, full_predictions)
        else:
            logits = self._logits(x, full_predictions)
        sel = self._topk(logits, cand, k)
        if isinstance(cand, SlabMap):
            up_map = cand.select(sel)
        else:
            up_map = ops.CoordMap(ops.take_keys(cand.keys, sel), cand.ts, True, cand.n_batches)  # predictions.C[mask] (:164)
        x = ST(on_map(up_map, gather(x.F, sel)), up_map)                        # _prune_coords (:101-128)
        if not self.dense and self.condition_ablation is None:
            x = seq3(x, self.conv_2)
        return x, ST(logits, cand), up_map


class ConditionEncoder(nn.Module):
    """model/blocks.py:185-251: Q-map pyramid -> (beta, gamma) per stage."""

    def __init__(self, C_in, N_scales, N_features, condition_ablation=None):
        super().__init__()
        self.num_stages = len(N_scales)
        self.condition_ablation = condition_ablation
        self.pre_conv = nn.Sequential(_conv(C_in, N_features[0]), ME.MinkowskiReLU(inplace=False))
        self.conv_layers, self.predict_layers, self.down_layers = nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
        for i in range(self.num_stages):
            down = _conv(N_features[i], N_features[i + 1], stride=2)
            self.down_layers.append(down)
            self._register_layers(down, f"down_layers_{i}")
            cl = nn.Sequential(_conv(N_features[i + 1], N_features[i + 1]), ME.MinkowskiReLU(inplace=False),
                               _conv(N_features[i + 1], N_features[i + 1]))
            self.conv_layers.append(cl)  # never called (:241) but part of the state dict
            self._register_layers(cl, f"conv_layers_{i}")
            pl = nn.Sequential(_conv(N_features[i + 1], N_scales[i]), ME.MinkowskiReLU(inplace=False),
                               _conv(N_scales[i], N_scales[i], k=1), ME.MinkowskiReLU(inplace=False),
                               _conv(N_scales[i], N_scales[i] * 2))
            self.predict_layers.append(pl)
            self._register_layers(pl, f"predict_layers_{i}")

    def _register_layers(self, layer, name):
        # the reference registers every parameter a second time under a flattened alias (:228-231)
        for pid, param in layer.named_parameters():
            self.register_parameter(f"{name}_{pid}".replace(".", "_"), param)

    def forward(self, Q):
        Q = apply_conv(Q, self.pre_conv[0], act=RELU)
        beta_gammas = []
        for i in range(self.num_stages):
            Q = apply_conv(Q, self.down_layers[i], out_map=Q.m.stride2())
            if self.condition_ablation is None:
                beta_gammas.append(seq5(Q, self.predict_layers[i]))
            elif self.condition_ablation == "condition_ablation":
                beta_gammas.append(Q)
        return Q, beta_gammas


def _count_per_batch(m):
    """AnalysisTransform.count_per_batch (model/transforms.py:65-71); maps are batch-sorted."""
    if isinstance(m, SlabMap):
        return [m.count_total()]
    if m.n == 0:
        return []
    nb = m.n_batches if m.n_batches is not None else int((m.keys.max() >> 48).item()) + 1
    if nb == 1:
        return [m.n]
    return torch.bincount((m.keys >> 48), minlength=nb).tolist()


class AnalysisTransform(nn.Module):
    """model/transforms.py:8-128."""

    def __init__(self, config):
        super().__init__()
        C_in, N1, N2, N3 = config["C_in"], config["N1"], config["N2"], config["N3"]
        self.condition_ablation = config["condition_ablation"] if "condition_ablation" in config else None
        if config["source_condition"]:
            self.cond_conv = nn.Sequential(_conv(C_in, 2), ME.MinkowskiReLU(inplace=False), _conv(2, 2))
        else:
            self.cond_conv = None
        self.pre_conv = nn.Sequential(_conv(C_in, N1), ME.MinkowskiReLU(inplace=False))
        self.down_1 = _conv(N1, N2, stride=2)
        self.down_2 = _conv(N2, N3, stride=2)
        self.down_3 = _conv(N3, N3, stride=2)
        self.scale_1 = ScaledBlock(N2, encode=True, scale=True)
        self.scale_2 = ScaledBlock(N3, encode=True, scale=True)
        self.scale_3 = ScaledBlock(N3, encode=True, scale=True)
        self.post_conv = _conv(N3, N3)
        self.condition_encoder = ConditionEncoder(C_in=2, N_scales=[N2, N2, N3], N_features=[2, 2, 2, 2],
                                                  condition_ablation=self.condition_ablation)

    def forward(self, x, Q):
        k = [_count_per_batch(x.m)]
        if self.cond_conv is not None:
            Q_plus = seq3(x, self.cond_conv)
            # Q.F + Q_plus.features_at_coordinates(Q.C)  (:94-98)
            Q = ST(on_map(Q.m, gather(Q_plus.F, rows_of(x.m, Q.m), add=Q.F)), Q.m)
        Q, beta_gammas = self.condition_encoder(Q)
        x = apply_conv(x, self.pre_conv[0], act=RELU)
        x = apply_conv(x, self.down_1, out_map=x.m.stride2())
        x = self.scale_1(x, beta_gammas[0])
        k.append(_count_per_batch(x.m))
        x = apply_conv(x, self.down_2, out_map=x.m.stride2())
        x = self.scale_2(x, beta_gammas[1])
        k.append(_count_per_batch(x.m))
        x = apply_conv(x, self.down_3, out_map=x.m.stride2())
        x = self.scale_3(x, beta_gammas[2])
        x = apply_conv(x, self.post_conv)
        Q = ST(on_map(x.m, gather(Q.F, rows_of(Q.m, x.m))), x.m)  # (:123-125)
        k.reverse()
        return x, Q, k


class SparseSynthesisTransform(nn.Module):
    """model/transforms.py:134-304."""

    def __init__(self, config):
        super().__init__()
        C_out, N1, N2, N3 = config["C_out"], config["N1"], config["N2"], config["N3"]
        dense = config["dense"] if "dense" in config.keys() else True
        if config["source_condition"]:
            self.cond_conv = nn.Sequential(_conv(N1, N1 // 2), ME.MinkowskiReLU(inplace=False), _conv(N1 // 2, 2))
        else:
            self.cond_conv = None
        self.pre_conv = nn.Sequential(_conv(N1, N1), ME.MinkowskiReLU(inplace=False))
        self.up_1 = GenerativeUpBlock(N1, N1, predict=True, dense=dense)
        self.up_2 = GenerativeUpBlock(N1, N2, predict=True, dense=dense)
        self.up_3 = GenerativeUpBlock(N2, N3, predict=True, dense=dense)
        self.scale_1 = ScaledBlock(N1, encode=False, scale=True)
        self.scale_2 = ScaledBlock(N1, encode=False, scale=True)
        self.scale_3 = ScaledBlock(N2, encode=False, scale=True)
        R = lambda: ME.MinkowskiReLU(inplace=False)
        self.post_conv = nn.Sequential(_conv(N3, N3), R(), _conv(N3, N3 // 2), R(), _conv(N3 // 2, C_out))
        self.q_pre_conv = nn.Sequential(_conv(2, 16), R(), _conv(16, 16, k=1), R(), _conv(16, 2))
        self.condition_ablation = config["condition_ablation"] if "condition_ablation" in config else None
        self.q_up_1 = GenerativeUpBlock(2, 2, condition_ablation=self.condition_ablation)
        self.q_up_2 = GenerativeUpBlock(2, 2, condition_ablation=self.condition_ablation)
        self.q_up_3 = GenerativeUpBlock(2, 2, condition_ablation=self.condition_ablation)
        self.q_predict_1 = nn.Sequential(_conv(2, N1), R(), _conv(N1, N1), R(), _conv(N1, N1 * 2))
        self.q_predict_2 = nn.Sequential(_conv(2, N1), R(), _conv(N1, N1), R(), _conv(N1, N1 * 2))
        self.q_predict_3 = nn.Sequential(_conv(2, N2), R(), _conv(N2, N2), R(), _conv(N2, N2 * 2))
        self.down_conv = ME.MinkowskiConvolution(in_channels=1, out_channels=1, kernel_size=3, stride=2, dimension=3)

    def forward(self, x, Q, coords=None, k=None, full_predictions=False):
        with stage("g_s.pre"):
            if self.cond_conv is not None:
                Q_plus = seq3(x, self.cond_conv)
                Q = ST(on_map(Q.m, gather(Q_plus.F, rows_of(x.m, Q.m), add=Q.F)), Q.m)  # (:258-265)
            x = apply_conv(x, self.pre_conv[0], act=RELU)
            Q = seq5(Q, self.q_pre_conv)
        predictions = []
        for s in (1, 2, 3):
            with stage(f"g_s.stage{s}.q_predict+scale"):
                # (beta, gamma) is read only at x's coordinates (model/blocks.py:37), so the last
                # q_predict convolution is evaluated there; at stage 1 Q still lives on h_q's dilated
                # map (132 k voxels for 18 k of x), afterwards Q and x share the map
                qp = getattr(self, f"q_predict_{s}")
                t = apply_conv(apply_conv(Q, qp[0], act=RELU), qp[2], act=RELU)
                beta_gamma = apply_conv(t, qp[4], out_map=x.m)
                x = getattr(self, f"scale_{s}")(x, beta_gamma)
            with stage(f"g_s.stage{s}.up"):
                x, pred, up_map = getattr(self, f"up_{s}")(x, k=k[s - 1], full_predictions=full_predictions)
                Q = getattr(self, f"q_up_{s}")(Q, up_map)
                Q = ST(Q.F, x.m)  # same coordinates after pruning: share the map object with x
            predictions.append(pred)
        with stage("g_s.post"):
            x = seq5(x, self.post_conv)
        if coords is not None:
            p1 = coords.m.stride2()
            p2 = p1.stride2()
            return x, [p2, p1, coords.m], predictions
        return x


class MeanScaleHyperprior_Map(CompressionModel):
    """model/entropy_models.py:253-414."""

    def __init__(self, config):
        super().__init__()
        Cb, Ch, CQ = config["C_bottleneck"], config["C_hyper_bottleneck"], config["C_Q"]
        self.entropy_bottleneck = EntropyBottleneck(Ch)
        self.gaussian_conditional = GaussianConditional(None)
        L = lambda: ME.MinkowskiLeakyReLU(inplace=False)
        R = lambda: ME.MinkowskiReLU(inplace=False)
        GT = lambda ci, co: ME.MinkowskiGenerativeConvolutionTranspose(
            in_channels=ci, out_channels=co, kernel_size=2, stride=2, bias=True, dimension=3)
        CT = lambda ci, co: ME.MinkowskiConvolutionTranspose(
            in_channels=ci, out_channels=co, kernel_size=3, stride=2, bias=True, dimension=3)
        self.h_a = nn.Sequential(_conv(Cb, Ch, bias=False), L(), _conv(Ch, Ch, stride=2, bias=False),
                                 _conv(Ch, Ch, bias=False), L(), _conv(Ch, Ch, stride=2, bias=False),
                                 _conv(Ch, Ch, bias=False))
        # the Sorted* shims of the reference (:12-102) only force canonical row order, which every
        # map of this implementation already has
        self.h_s = nn.Sequential(_conv(Ch, Ch), GT(Ch, Ch), L(), _conv(Ch, Ch), GT(Ch, Cb * 3 // 2), L(),
                                 _conv(Cb * 3 // 2, Cb * 2))
        self.h_q = nn.Sequential(_conv(Ch, Ch), CT(Ch, Ch), R(), _conv(Ch, Ch), CT(Ch, Ch), R(), _conv(Ch, CQ))

    def run_h_a(self, y):
        z = apply_conv(y, self.h_a[0], act=LEAKY)
        z = apply_conv(z, self.h_a[2], out_map=z.m.stride2())
        z = apply_conv(z, self.h_a[3], act=LEAKY)
        z = apply_conv(z, self.h_a[5], out_map=z.m.stride2())
        return apply_conv(z, self.h_a[6])

    def run_h_s(self, z_hat, at_map=None):
        """h_s; with at_map the last convolution is evaluated only at those coordinates
        (gaussian_params.features_at_coordinates(y.C), :364,401)."""
        g = apply_conv(z_hat, self.h_s[0])
        g = apply_conv(g, self.h_s[1], out_map=g.m.expand(2), transpose=True, act=LEAKY)
        g = apply_conv(g, self.h_s[3])
        g = apply_conv(g, self.h_s[4], out_map=g.m.expand(2), transpose=True, act=LEAKY)
        return apply_conv(g, self.h_s[6], out_map=at_map)

    def run_h_q(self, z_hat):
        """h_q: z_hat has a fresh coordinate manager in the reference (:356-359,394-397), so both
        ConvolutionTranspose layers generate the 27-dilated map."""
        q = apply_conv(z_hat, self.h_q[0])
        q = apply_conv(q, self.h_q[1], out_map=q.m.expand(3), transpose=True, act=RELU)
        q = apply_conv(q, self.h_q[3])
        q = apply_conv(q, self.h_q[4], out_map=q.m.expand(3), transpose=True, act=RELU)
        return apply_conv(q, self.h_q[6])

    def forward(self, y):
        """Forward with likelihoods (:309-337); in training mode with CompressAI's additive-noise
        quantisation and differentiable likelihoods."""
        if self.training:
            return self._forward_train(y)
        z = self.run_h_a(y)
        med, packed = self.entropy_bottleneck._device_params(z.F.device)
        z_hat_f, z_lik = ops.eb_likelihood(z.F, med, packed, 1e-9)
        z_hat = ST(z_hat_f, z.m)
        params = self.run_h_s(z_hat, at_map=y.m)
        Q_hat = self.run_h_q(z_hat)
        gc = self.gaussian_conditional
        y_hat_f, y_lik = ops.gc_likelihood(y.F, params.F, None, gc._scale_bound, 1e-9)
        return ST(y_hat_f, y.m), Q_hat, (y_lik.unsqueeze(0), z_lik.unsqueeze(0))

    def _forward_train(self, y):
        z = self.run_h_a(y)
        z_hat_cn, z_lik = self.entropy_bottleneck(z.F.t().unsqueeze(0), training=True)  # (:313)
        z_hat = ST(z_hat_cn[0].t().contiguous(), z.m)
        params = self.run_h_s(z_hat, at_map=y.m)       # gaussian_params at y.C (:324-327)
        Q_hat = self.run_h_q(z_hat)
        gc = self.gaussian_conditional
        y_hat_f, y_lik = gc.forward_rows(y.F, params.F, training=True)  # (:330)
        return ST(y_hat_f, y.m), Q_hat, (y_lik.t().unsqueeze(0), z_lik)

    def compress(self, y):
        return drive(self.compress_gen(y))

    def decompress(self, points, strings, shape):
        return drive(self.decompress_gen(points, strings, shape))

    def compress_gen(self, y):
        """(:341-381).  y is canonical already, so sort_tensor (:346-347) is the identity.
        Generator: yields compressai_compat.Wait objects where the CPU coder is involved."""
        with stage("enc.h_a"):
            z = self.run_h_a(y)
        shape = [z.m.n]
        with stage("enc.z_coder"):
            z_sym, z_hat_f = self.entropy_bottleneck.quantize_rows(z.F)
            z_string = self.entropy_bottleneck.compress_symbols(z_sym)
        z_hat = ST(z_hat_f, z.m)  # == decompress(compress(z)) (:352-353): the coder is lossless
        with stage("enc.h_s"):
            params = self.run_h_s(z_hat, at_map=y.m)
        if STAGE_MS is not None:
            with stage("enc.y_coder"):
                y_string, y_sym, y_idx = self.gaussian_conditional.compress_fused(y.F, params.F, None)
        else:
            y_string, y_sym, y_idx = yield from self.gaussian_conditional.compress_fused_gen(y.F, params.F, None)
        self.last = {"z": z, "z_hat": z_hat, "params": params, "y_symbols": y_sym, "indexes": y_idx,
                     "z_symbols": z_sym}
        return [y.m, z.m], [[y_string], [z_string]], shape

    def decompress_gen(self, points, strings, shape):
        """(:384-414).  points = [y map, z map], both canonical (sort_points, :389-390)."""
        y_map, z_map = points
        n_z = int(shape[0]) if isinstance(shape, (list, tuple)) else int(shape)
        if n_z != z_map.n:  # a truncated / tampered header must not size tensors the kernels index by the map
            raise PccError(f"corrupt stream: header says {n_z} hyper-latent rows, the latent coordinates give {z_map.n}")
        with stage("dec.z_coder"):
            z_hat_cn = self.entropy_bottleneck.decompress(strings[1], shape)
            z_hat = ST(z_hat_cn[0].t().contiguous(), z_map)
        with stage("dec.h_q+h_s"):
            Q_hat = self.run_h_q(z_hat)
            params = self.run_h_s(z_hat, at_map=y_map)
        if STAGE_MS is not None:
            with stage("dec.y_coder"):
                y_hat_f = self.gaussian_conditional.decompress_fused(strings[0][0], params.F, None, y_map.n,
                                                                     params.F.shape[1] // 2)
        else:
            y_hat_f = yield from self.gaussian_conditional.decompress_fused_gen(
                strings[0][0], params.F, None, y_map.n, params.F.shape[1] // 2)
        self.last = {"z_hat": z_hat, "params": params}
        return ST(y_hat_f, y_map), Q_hat


class MeanScaleHyperprior(MeanScaleHyperprior_Map):
    """model/entropy_models.py:104-250: the variant without h_q, selected only when the config has
    an ``entropy_model_map`` section (no shipped config does)."""

    def __init__(self, config):
        cfg = dict(config)
        cfg.setdefault("C_Q", 2)
        super().__init__(cfg)
        del self.h_q

    def run_h_q(self, z_hat):
        return None


def _canonical(coords, feats, ts, share_with=None):
    """Sort (coords, feats) into canonical order.  Reuses `share_with`'s map object when the
    coordinate sets are identical, so both branches share kernel maps."""
    m, perm = ops.canonical_map(coords, ts)
    f = ops.gather_rows(feats.float().contiguous(), perm.to(torch.int32))
    if share_with is not None and share_with.n == m.n and torch.equal(share_with.keys, m.keys):
        m = share_with
    return ST(f, m)


class ColorModel(CompressionModel):
    """model/model.py:15-395 (encode/decode path; the G-PCC side channel and the file container,
    :214-395, are outside the hot path — SURVEY.md §8f)."""

    def __init__(self, config):
        super().__init__()
        self.g_a = AnalysisTransform(config["g_a"])
        self.g_s = SparseSynthesisTransform(config["g_s"])
        if "entropy_model_map" in config.keys():
            self.entropy_model = MeanScaleHyperprior(config["entropy_model"])
            self.entropy_model_map = MeanScaleHyperprior(config["entropy_model_map"])
        else:
            self.entropy_model = MeanScaleHyperprior_Map(config["entropy_model"])
            self.entropy_model_map = None
        self.stage_ms = {}

    def update(self):
        self.entropy_model.update(force=True)
        if self.entropy_model_map is not None:
            self.entropy_model_map.update(force=True)

    def aux_loss(self):
        if self.entropy_model_map is None:
            return self.entropy_model.aux_loss()
        return self.entropy_model.aux_loss() + self.entropy_model_map.aux_loss()

    @property
    def device(self):
        return self.g_s.down_conv.kernel.device  # model/model.py:179

    def _inputs(self, x, Q):
        """model/model.py:114-123 + canonical ordering.  x [N,6] (xyz, rgb); Q: SparseTensor-like
        with .C [N,4] and .F [N,2] (utils.py:436-445), or a ([N,4], [N,2]) tuple."""
        dev = self.device
        x = x.to(dev, non_blocking=True)
        N = x.shape[0]
        if N == 0 or x.dim() != 2 or x.shape[1] < 6:
            # the reference hands this to ME.SparseTensor / torch.cat, which reject it with a shape error
            raise PccError(f"compress needs a non-empty [N, 6] cloud (xyz, rgb); got {tuple(x.shape)}")
        pts = torch.cat([torch.zeros((N, 1), device=dev), x[:, :3]], dim=1).int()
        feats = torch.cat([torch.ones((N, 1), device=dev), x[:, 3:6]], dim=1)
        inp = _canonical(pts, feats, 1)
        qc, qf = (Q if isinstance(Q, (tuple, list)) else (Q.C, Q.F))
        qc = qc.to(dev, non_blocking=True)
        if qc.is_floating_point():
            qc = torch.floor(qc).int()
        Qt = _canonical(qc, qf.to(dev, non_blocking=True), 1, share_with=inp.m)
        return inp, Qt

    def _inputs_sparse(self, x, Q):
        """model/model.py:58-66 for a (batched) SparseTensor-like x with .C [N,4] and .F [N,3]."""
        dev = self.device
        xc, xf = x.C.to(dev), x.F.to(dev)
        feats = torch.cat([torch.ones((xc.shape[0], 1), device=dev), xf], dim=1)
        inp = _canonical(xc.int(), feats, 1)
        qc, qf = (Q if isinstance(Q, (tuple, list)) else (Q.C, Q.F))
        qc = qc.to(dev)
        if qc.is_floating_point():
            qc = torch.floor(qc).int()
        return inp, _canonical(qc, qf.to(dev), 1, share_with=inp.m)

    def forward(self, x, Q, Lambda=None):
        """model/model.py:51-93: likelihoods instead of strings.  x: [N,6] tensor (one cloud), a
        SparseTensor-like (.C with batch ids, .F colours: train.py:176-183) or an internal ST.  In
        training mode (``model.train()``) the pass is recorded for ``loss.backward()``; otherwise it
        runs without a tape on the fused kernels."""
        with torch.set_grad_enabled(self.training and torch.is_grad_enabled()):
            if isinstance(x, ST):
                inp, Qt = x, Q
            elif torch.is_tensor(x):
                inp, Qt = self._inputs(x, Q)
            else:
                inp, Qt = self._inputs_sparse(x, Q)
            y, Qy, k = self.g_a(inp, Qt)
            if self.entropy_model_map is None:
                y_hat, Q_hat, lik = self.entropy_model(y)
                likelihoods = {"y": lik[0], "z": lik[1]}
            else:  # (:78-81)
                y_hat, _, y_lik = self.entropy_model(y)
                Q_hat, _, Q_lik = self.entropy_model_map(Qy)
                likelihoods = {"y": [y_lik[0], Q_lik[0]], "z": [y_lik[1], Q_lik[1]]}
            x_hat, points, predictions = self.g_s(y_hat, Q_hat, coords=ST(None, inp.m), k=k,
                                                  full_predictions=not self.training)
            return {"prediction": x_hat, "points": points, "occ_predictions": predictions, "q_map": Lambda,
                    "likelihoods": likelihoods, "input": inp}

    @torch.no_grad()
    def compress(self, x, Q, path=None):
        """model/model.py:95-147: returns (strings, shape, k, coordinates), or writes `path`."""
        return drive(self.compress_gen(x, Q, path))

    @torch.no_grad()
    def decompress(self, path=None, coordinates=None, strings=None, shape=None, k=None):
        """model/model.py:152-208 -> [N, 6] float (xyz, rgb quantised to 8 bit)."""
        return drive(self.decompress_gen(path, coordinates, strings, shape, k))

    def save_bitstream(self, path, points, strings, shape, k):
        """model/model.py:214-263."""
        from .bitstream import save_bitstream
        return save_bitstream(path, points, strings, shape, k)

    def load_bitstream(self, path):
        """model/model.py:267-315."""
        from .bitstream import load_bitstream
        return load_bitstream(path, self.device)

    def gpcc_encode(self, points, directory=None):
        """model/model.py:318-358 (own octree code instead of the tmc3 subprocess; `directory` unused)."""
        from .bitstream import gpcc_encode
        return gpcc_encode(points)

    def gpcc_decode(self, bin, directory=None):
        """model/model.py:360-395."""
        from .bitstream import gpcc_decode
        return gpcc_decode(bin, self.device)

    def compress_gen(self, x, Q, path=None):
        """Generator form of compress (yields compressai_compat.Wait objects at the coder)."""
        with stage("enc.inputs"):
            inp, Qt = self._inputs(x, Q)
        with stage("enc.g_a"):
            y, Qy, k = self.g_a(inp, Qt)
        if self.entropy_model_map is None:
            points, strings, shape = yield from self.entropy_model.compress_gen(y)
        else:
            points, y_strings, y_shape = yield from self.entropy_model.compress_gen(y)
            _, Q_strings, Q_shape = yield from self.entropy_model_map.compress_gen(Qy)
            strings, shape = [y_strings, Q_strings], [y_shape, Q_shape]
        self.last = {"y": y, "Q": Qy}
        coordinates = y.m.coords
        if path:
            from .bitstream import save_bitstream
            save_bitstream(path, coordinates, strings, shape, k)
            return None
        return strings, shape, k, coordinates

    def decompress_gen(self, path=None, coordinates=None, strings=None, shape=None, k=None):
        """Generator form of decompress."""
        dev = self.device
        if path:
            from .bitstream import load_bitstream
            coordinates, strings, shape, k = load_bitstream(path, dev)
        coordinates = coordinates.to(dev)
        if coordinates.shape[1] == 3:
            coordinates = torch.cat([torch.zeros((coordinates.shape[0], 1), device=dev,
                                                 dtype=coordinates.dtype), coordinates], dim=1)
        if coordinates.is_floating_point():
            coordinates = torch.floor(coordinates)
        if coordinates.shape[0] == 0 or len(k) != 3 or any(len(ki) == 0 for ki in k):
            raise PccError("decompress needs the latent coordinates and the three per-batch point counts of a "
                           f"non-empty cloud; got {tuple(coordinates.shape)} coordinates, k = {k}")
        with stage("dec.maps"):
            y_map, _ = ops.canonical_map(coordinates.int(), 8)
            z_map = y_map.stride2().stride2()  # g_s.down_conv twice, only .C is used (:188-191)
        if self.entropy_model_map is None:
            y_hat, Q_hat = yield from self.entropy_model.decompress_gen([y_map, z_map], strings, shape)
        else:
            y_hat, _ = yield from self.entropy_model.decompress_gen([y_map, z_map], strings[0], shape[0])
            Q_hat, _ = yield from self.entropy_model_map.decompress_gen([y_map, z_map], strings[1], shape[1])
        self.last = {"y_hat": y_hat, "Q_hat": Q_hat}
        x_hat = self.g_s(y_hat, Q_hat, k=k)
        self.last["x_hat"] = x_hat
        with stage("dec.finalize"):
            return ops.color_finalize(x_hat.m.keys, x_hat.F)  # (:206-207)


# ------------------------------------------------------------------------------------------
# Synthetic weights (no checkpoint is published: README.md:218).  ME's default init shrinks
# activations by ~0.15x per layer, which after ~40 layers leaves every logit equal to its bias;
# this variance-preserving init (gain tuned on the oracle) keeps activations O(1) so top-k and the quantiser see non-degenerate data.
def _effective_taps(m):
    if m.kernel_volume == 1:
        return 1.0
    if isinstance(m, (ME.MinkowskiGenerativeConvolutionTranspose, ME.MinkowskiConvolutionTranspose)):
        return 1.0 if m.kernel_size == 2 else 3.4
    return 13.0 if m.stride == 1 else 8.0


SYNTHETIC_GAIN = 1.1  # tuned with the oracle so that y, the scale indexes and the logits are non-degenerate


def init_synthetic(model, seed=0):
    """Deterministic (numpy PCG64) weights for every convolution; entropy-bottleneck parameters
    keep CompressAI's initial values except the biases, which are re-drawn from the same stream."""
    import numpy as np
    rng = np.random.default_rng(seed)
    seen = set()
    with torch.no_grad():
        for name, m in model.named_modules():
            if isinstance(m, ME._ConvBase):
                if id(m.kernel) in seen:
                    continue
                seen.add(id(m.kernel))
                s = SYNTHETIC_GAIN * float(np.sqrt(3.0 / (m.in_channels * _effective_taps(m))))
                m.kernel.copy_(torch.from_numpy(rng.uniform(-s, s, size=tuple(m.kernel.shape)).astype(np.float32)))
                if m.bias is not None:
                    m.bias.copy_(torch.from_numpy(rng.uniform(-0.1, 0.1, size=tuple(m.bias.shape)).astype(np.float32)))
            elif isinstance(m, EntropyBottleneck):
                for i in range(5):
                    b = getattr(m, f"_bias{i}")
                    b.copy_(torch.from_numpy(rng.uniform(-0.5, 0.5, size=tuple(b.shape)).astype(np.float32)))
    return model


def export_params(model):
    """state_dict -> {name: numpy array} for the CPU oracle (tests / bench baseline only)."""
    return {k: v.detach().cpu().numpy() for k, v in model.state_dict().items()}


OURS_CONFIG = {  # configs/Ours.yaml:6-23
    "entropy_model": {"type": "MeanScaleHyperprior_map", "C_bottleneck": 128, "C_hyper_bottleneck": 128, "C_Q": 2},
    "g_a": {"C_in": 4, "N1": 64, "N2": 128, "N3": 128, "source_condition": True},
    "g_s": {"C_out": 3, "N1": 128, "N2": 128, "N3": 64, "source_condition": True},
}
ABLATION_CONFIG = {  # configs/Ablation_NoCondition_Convolution.yaml:6-25
    "entropy_model": {"type": "MeanScaleHyperprior_map", "C_bottleneck": 128, "C_hyper_bottleneck": 128, "C_Q": 2},
    "g_a": {"C_in": 4, "N1": 64, "N2": 128, "N3": 128, "source_condition": False,
            "condition_ablation": "condition_ablation"},
    "g_s": {"C_out": 3, "N1": 128, "N2": 128, "N3": 64, "source_condition": False,
            "condition_ablation": "condition_ablation"},
}


def load_config(path):
    """The ``model:`` section of a reference YAML (train.py:113-118)."""
    import yaml
    with open(path) as f:
        return yaml.safe_load(f)["model"]
