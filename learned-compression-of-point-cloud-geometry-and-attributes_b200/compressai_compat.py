"""CompressAI-shaped entropy modules over the B200 kernels and the C++ range coder.

Exports the names the reference imports (model/entropy_models.py:9-10, model/model.py:4):
``EntropyBottleneck``, ``GaussianConditional``, ``CompressionModel`` — same constructor
arguments, parameter / buffer names (``quantiles``, ``_matrix{i}``, ``_bias{i}``, ``_factor{i}``,
``_quantized_cdf``, ``_cdf_length``, ``_offset``, ``scale_table`` …) and call signatures
(tensors shaped [1, C, N]).  Semantics: SURVEY.md Appendix B.

Device work (quantise, CDF index, likelihood) runs in pcc_* kernels; the range coder is the
reference's sequential CPU algorithm (csrc/host_coder.cpp) and its time is accumulated in
``coder_stats`` so callers can report it separately from the GPU path.
"""
import math
import threading
import time

import numpy as np
import torch
import torch.nn as nn

from . import ops
from ._lib import PccError

coder_stats = {"encode_s": 0.0, "decode_s": 0.0, "d2h_s": 0.0, "h2d_s": 0.0, "symbols": 0,
               "d2h_bytes": 0, "h2d_bytes": 0}


_stats_lock = threading.Lock()


def _stat(key, value):
    """coder_stats is updated from the launch and coder threads of every in-flight frame."""
    with _stats_lock:
        coder_stats[key] += value


def reset_coder_stats():
    with _stats_lock:
        for k in coder_stats:
            coder_stats[k] = 0 if isinstance(coder_stats[k], int) else 0.0


def get_scale_table(min=0.11, max=256, levels=64):
    """compressai.models.utils.get_scale_table."""
    return torch.exp(torch.linspace(math.log(min), math.log(max), levels))


class _LowerBoundFn(torch.autograd.Function):
    """compressai.ops.LowerBoundFunction: max(x, bound); the gradient passes where x >= bound or
    where it would move x up (grad < 0)."""

    @staticmethod
    def forward(ctx, x, bound):
        ctx.save_for_backward(x, bound)
        return torch.max(x, bound)

    @staticmethod
    def backward(ctx, g):
        x, bound = ctx.saved_tensors
        return ((x >= bound) | (g < 0)).to(g.dtype) * g, None


class LowerBound(nn.Module):
    """compressai.ops.LowerBound: max(x, bound) with the pass-through gradient rule; carries the
    ``bound`` buffer."""

    def __init__(self, bound):
        super().__init__()
        self.register_buffer("bound", torch.Tensor([float(bound)]))

    def forward(self, x):
        return _LowerBoundFn.apply(x, self.bound.to(x.device))


# Additive quantisation noise of the training path (compressai EntropyModel.quantize, mode "noise":
# inputs + U(-0.5, 0.5)).  Tests replace the source to feed the oracle and the product the same draw.
_noise_source = None


def set_noise_source(fn):
    """fn(tensor) -> noise tensor of the same shape/device in [-0.5, 0.5); None restores torch's RNG."""
    global _noise_source
    _noise_source = fn


def quantization_noise(like):
    if _noise_source is not None:
        return _noise_source(like).to(like.device, like.dtype)
    return torch.empty_like(like).uniform_(-0.5, 0.5)


# ------------------------------------------------------------------------------ asynchronous coder
# The range coder is the reference's sequential CPU algorithm; it runs in a small thread pool (the
# ctypes call releases the GIL) so that a scheduler (pipeline.FramePipeline) can keep other frames'
# kernels flowing while one frame is being coded.  Code that needs the coder is written as a
# generator yielding `Wait` objects; `drive()` runs such a generator synchronously.
import concurrent.futures as _cf

_coder_pool = _cf.ThreadPoolExecutor(max_workers=4, thread_name_prefix="pcc-coder")


class Wait:
    """What a codec generator is waiting for: a CUDA event (device work / copy) or a coder future."""
    __slots__ = ("event", "future")

    def __init__(self, event=None, future=None):
        self.event, self.future = event, future

    def ready(self):
        if self.event is not None and not self.event.query():
            return False
        return self.future is None or self.future.done()

    def block(self):
        if self.event is not None:
            self.event.synchronize()
        if self.future is not None:
            self.future.result()


def drive(gen):
    """Run a codec generator to completion, blocking at every wait point."""
    try:
        w = next(gen)
        while True:
            w.block()
            w = gen.send(None)
    except StopIteration as e:
        return e.value


_pinned_free = {}
_pinned_lock = threading.Lock()  # launch threads and coder threads of several in-flight frames share the pool


def _pinned(shape, dtype):
    """Page-locked host tensor from a free-list (cudaHostAlloc is far too slow for the hot path)."""
    key = (tuple(shape), dtype)
    with _pinned_lock:
        lst = _pinned_free.get(key)
        hit = lst.pop() if lst else None
    if hit is not None:
        t, ev = hit
        if ev is not None:
            ev.synchronize()
        return t
    return torch.empty(shape, dtype=dtype, pin_memory=True)


def _pinned_release(t, after_event=None):
    with _pinned_lock:
        _pinned_free.setdefault((tuple(t.shape), t.dtype), []).append((t, after_event))


def _d2h_async(*tensors):
    """Device tensors -> pinned host tensors (async on the current stream) + the event to wait on."""
    outs = []
    for t in tensors:
        h = _pinned(t.shape, t.dtype)
        h.copy_(t, non_blocking=True)
        _stat("d2h_bytes", h.numel() * h.element_size())
        outs.append(h)
    ev = torch.cuda.Event()
    ev.record()
    return outs, ev


def _h2d_async(a, device):
    """numpy array -> device tensor through a pinned staging buffer (async on the current stream)."""
    h = _pinned(a.shape, torch.from_numpy(a).dtype)
    h.numpy()[...] = a
    t = h.to(device, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record()
    _pinned_release(h, ev)
    _stat("h2d_bytes", a.nbytes)
    return t


def _rows_cn(x):
    """[1, C, N] (typically ``F.t().unsqueeze(0)``) -> contiguous row-major [N, C] without a copy
    when the input is a transposed view."""
    if x.dim() != 3 or x.shape[0] != 1:
        raise PccError("entropy modules take [1, C, N] tensors (batch of one, as the reference)")
    t = x[0].t()
    return t if t.is_contiguous() else t.contiguous()


def _to_host(t):
    t0 = time.perf_counter()
    a = t.cpu().numpy()
    _stat("d2h_s", time.perf_counter() - t0)
    _stat("d2h_bytes", a.nbytes)
    return a


def _to_dev(a, device):
    t0 = time.perf_counter()
    t = torch.from_numpy(a).to(device)
    _stat("h2d_s", time.perf_counter() - t0)
    _stat("h2d_bytes", a.nbytes)
    return t


class EntropyModel(nn.Module):
    def __init__(self, likelihood_bound=1e-9, entropy_coder=None, entropy_coder_precision=16):
        super().__init__()
        self.entropy_coder_precision = int(entropy_coder_precision)
        self.use_likelihood_bound = likelihood_bound > 0
        self._likelihood_bound = float(likelihood_bound)
        if self.use_likelihood_bound:
            self.likelihood_lower_bound = LowerBound(likelihood_bound)
        self.register_buffer("_offset", torch.IntTensor())
        self.register_buffer("_quantized_cdf", torch.IntTensor())
        self.register_buffer("_cdf_length", torch.IntTensor())
        self._host_tables = None
        self._dev_cache = {}

    _resizable_buffers = ("_offset", "_quantized_cdf", "_cdf_length")

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                              error_msgs):
        """CompressAI's ``update_registered_buffers`` (compressai/models/utils.py), done where it cannot be
        missed: the coder tables are registered empty and get their size in ``update()``, so a checkpoint
        saved AFTER ``update()`` (the reference's: train.py:291 calls ``model.update()`` before
        ``state_dict()``, evaluate.py:81 loads that file) carries shapes this module does not have yet.
        The buffers are re-created with the incoming shapes before the default copy, whatever class the
        parent model is (this repo's ColorModel or the reference's)."""
        for name in self._resizable_buffers:
            src = state_dict.get(prefix + name)
            buf = getattr(self, name)
            if src is not None and tuple(src.shape) != tuple(buf.shape):
                setattr(self, name, torch.empty(tuple(src.shape), dtype=buf.dtype, device=buf.device))
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                                      error_msgs)
        self._host_tables = None  # rebuilt from the loaded buffers on first use
        self._dev_cache = {}

    def _set_tables(self, cdf, length, offset):
        self._host_tables = (np.ascontiguousarray(cdf, np.int32), np.ascontiguousarray(length, np.int32),
                             np.ascontiguousarray(offset, np.int32))
        dev = self._quantized_cdf.device
        self._quantized_cdf = torch.from_numpy(self._host_tables[0].copy()).to(dev)
        self._cdf_length = torch.from_numpy(self._host_tables[1].copy()).to(dev)
        self._offset = torch.from_numpy(self._host_tables[2].copy()).to(dev)
        self._dev_cache = {}

    def _tables(self):
        if self._host_tables is None:
            if self._quantized_cdf.numel() == 0:
                raise PccError("entropy model not initialised: call update() before compress()")
            self._host_tables = (self._quantized_cdf.cpu().numpy().astype(np.int32),
                                 self._cdf_length.cpu().numpy().astype(np.int32),
                                 self._offset.cpu().numpy().astype(np.int32))
        return self._host_tables

    def _encode_host(self, sym, idx):
        """numpy symbols/indexes -> bytes (runs in a coder thread)."""
        cdf, length, offset = self._tables()
        t0 = time.perf_counter()
        s = ops.rans_encode(sym, idx, cdf, length, offset)
        _stat("encode_s", time.perf_counter() - t0)
        _stat("symbols", sym.size)
        return s

    def _encode(self, symbols_dev, indexes_dev):
        sym, idx = _to_host(symbols_dev), _to_host(indexes_dev)
        cdf, length, offset = self._tables()
        t0 = time.perf_counter()
        s = ops.rans_encode(sym, idx, cdf, length, offset)
        _stat("encode_s", time.perf_counter() - t0)
        _stat("symbols", sym.size)
        return s

    def _decode(self, string, indexes_host):
        cdf, length, offset = self._tables()
        t0 = time.perf_counter()
        out = ops.rans_decode(string, indexes_host, cdf, length, offset)
        _stat("decode_s", time.perf_counter() - t0)
        _stat("symbols", out.size)
        return out


class EntropyBottleneck(EntropyModel):
    """compressai.entropy_models.EntropyBottleneck (SURVEY.md Appendix B1;
    model/entropy_models.py:269,313,352-353,393)."""

    def __init__(self, channels, *args, tail_mass=1e-9, init_scale=10, filters=(3, 3, 3, 3), **kwargs):
        super().__init__(*args, **kwargs)
        if tuple(filters) != (3, 3, 3, 3):
            raise PccError("only the default filters (3,3,3,3) are supported")
        self.channels = int(channels)
        self.filters = tuple(int(f) for f in filters)
        self.init_scale = float(init_scale)
        self.tail_mass = float(tail_mass)
        f = (1,) + self.filters + (1,)
        scale = self.init_scale ** (1 / (len(self.filters) + 1))
        for i in range(len(self.filters) + 1):
            init = np.log(np.expm1(1 / scale / f[i + 1]))
            matrix = torch.Tensor(channels, f[i + 1], f[i])
            matrix.data.fill_(init)
            self.register_parameter(f"_matrix{i:d}", nn.Parameter(matrix))
            bias = torch.Tensor(channels, f[i + 1], 1)
            nn.init.uniform_(bias, -0.5, 0.5)
            self.register_parameter(f"_bias{i:d}", nn.Parameter(bias))
            if i < len(self.filters):
                factor = torch.Tensor(channels, f[i + 1], 1)
                nn.init.zeros_(factor)
                self.register_parameter(f"_factor{i:d}", nn.Parameter(factor))
        self.quantiles = nn.Parameter(torch.Tensor(channels, 1, 3))
        init = torch.Tensor([-self.init_scale, 0, self.init_scale])
        self.quantiles.data = init.repeat(self.quantiles.size(0), 1, 1)
        target = np.log(2 / self.tail_mass - 1)
        self.register_buffer("target", torch.Tensor([-target, 0, target]))

    # -- host-side table construction (one-off; ColorModel.update, model/model.py:30-36)
    def update(self, force=False):
        if self._offset.numel() > 0 and not force:
            return False
        g = lambda name: getattr(self, name).detach().cpu().numpy().astype(np.float32)
        cdf, length, offset = ops.factorized_cdf_tables(
            [g(f"_matrix{i}") for i in range(5)], [g(f"_bias{i}") for i in range(5)],
            [g(f"_factor{i}") for i in range(4)], g("quantiles"))
        self._set_tables(cdf, length, offset)
        return True

    def _device_params(self, device):
        """medians [C] and the 58 pre-transformed floats per channel the likelihood kernel reads."""
        key = str(device)
        hit = self._dev_cache.get(key)
        if hit is not None:
            return hit
        with ops.cache_lock, torch.no_grad():
            parts = []
            for i in range(5):
                parts.append(torch.nn.functional.softplus(getattr(self, f"_matrix{i}").double()).reshape(self.channels, -1))
                parts.append(getattr(self, f"_bias{i}").double().reshape(self.channels, -1))
                if i < 4:
                    parts.append(torch.tanh(getattr(self, f"_factor{i}").double()).reshape(self.channels, -1))
            packed = torch.cat(parts, dim=1).float().contiguous().to(device)
            med = self.quantiles[:, 0, 1].detach().float().contiguous().to(device)
            ops.publish(med, packed)
            self._dev_cache[key] = (med, packed)
        return self._dev_cache[key]

    def _get_medians(self):
        return self.quantiles[:, :, 1:2].detach()

    def forward(self, x, training=None):
        if training if training is not None else self.training:
            return self._forward_train(x)
        z = _rows_cn(x)
        med, packed = self._device_params(z.device)
        z_hat, lik = ops.eb_likelihood(z, med, packed, self._likelihood_bound if self.use_likelihood_bound else 0.0)
        return z_hat.t().unsqueeze(0), lik.unsqueeze(0)

    def _forward_train(self, x):
        """EntropyBottleneck.forward(training=True): x [1, C, N] -> (x + noise, likelihood), both
        [1, C, N], on the autograd tape.  The factorised density is C x 58 parameters applied to
        C x N values (N = 1.2 k per 800 k-point frame): plain tensor expressions."""
        self._dev_cache.clear()  # parameters are about to change: eval-mode tables are stale
        values = x.permute(1, 0, 2).reshape(self.channels, 1, -1)
        outputs = values + quantization_noise(values)
        lower = self._logits_cumulative(outputs - 0.5)
        upper = self._logits_cumulative(outputs + 0.5)
        sign = -torch.sign(lower + upper).detach()
        lik = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))
        if self.use_likelihood_bound:
            lik = self.likelihood_lower_bound(lik)
        shape = (self.channels, x.shape[0], -1)
        return outputs.reshape(shape).permute(1, 0, 2), lik.reshape(shape).permute(1, 0, 2)

    def quantize_rows(self, z):
        """z [N, C] -> (symbols [C, N] int32 device, z_hat [N, C])."""
        med, _ = self._device_params(z.device)
        return ops.eb_quantize(z, med)

    def compress(self, x):
        z = _rows_cn(x)
        sym, _ = self.quantize_rows(z)
        return [self.compress_symbols(sym)]

    def compress_symbols(self, sym_dev):
        c, n = sym_dev.shape
        idx = np.repeat(np.arange(c, dtype=np.int32)[:, None], n, axis=1)
        sym = _to_host(sym_dev)
        cdf, length, offset = self._tables()
        t0 = time.perf_counter()
        s = ops.rans_encode(sym, idx, cdf, length, offset)
        _stat("encode_s", time.perf_counter() - t0)
        _stat("symbols", sym.size)
        return s

    def decompress(self, strings, size):
        n = int(size[0]) if isinstance(size, (list, tuple, torch.Size)) else int(size)
        device = self.quantiles.device
        idx = np.repeat(np.arange(self.channels, dtype=np.int32)[:, None], n, axis=1)
        sym = self._decode(strings[0], idx).reshape(self.channels, n)
        med, _ = self._device_params(device)
        z_hat = ops.eb_dequantize(_to_dev(sym, device), med)
        return z_hat.t().unsqueeze(0)

    def _logits_cumulative(self, inputs, stop_gradient=False):
        p = (lambda t: t.detach()) if stop_gradient else (lambda t: t)
        logits = inputs
        for i in range(5):
            logits = torch.matmul(torch.nn.functional.softplus(p(getattr(self, f"_matrix{i}"))), logits)
            logits = logits + p(getattr(self, f"_bias{i}"))
            if i < 4:
                logits = logits + torch.tanh(p(getattr(self, f"_factor{i}"))) * torch.tanh(logits)
        return logits

    def loss(self):
        """Auxiliary loss (model/model.py:38-47, train.py:199-203): only the quantiles learn from it
        (CompressAI evaluates the density with stop_gradient=True here)."""
        logits = self._logits_cumulative(self.quantiles, stop_gradient=True)
        return torch.abs(logits - self.target).sum()


class GaussianConditional(EntropyModel):
    """compressai.entropy_models.GaussianConditional (SURVEY.md Appendix B2;
    model/entropy_models.py:270,330,371-372,407-408)."""
    _resizable_buffers = EntropyModel._resizable_buffers + ("scale_table",)

    def __init__(self, scale_table, *args, scale_bound=0.11, tail_mass=1e-9, **kwargs):
        super().__init__(*args, **kwargs)
        if scale_table is not None and (len(scale_table) < 1 or len(scale_table) > 64):
            raise PccError("scale_table must have 1..64 entries")
        self.tail_mass = float(tail_mass)
        if scale_bound is None and scale_table:
            scale_bound = float(scale_table[0])
        self._scale_bound = float(scale_bound)
        self.lower_bound_scale = LowerBound(scale_bound)
        self.register_buffer("scale_table", torch.Tensor(tuple(float(s) for s in scale_table))
                             if scale_table is not None else torch.Tensor())
        self.register_buffer("scale_bound", torch.Tensor([float(scale_bound)]))

    def update_scale_table(self, scale_table, force=False):
        if self._offset.numel() > 0 and not force:
            return False
        dev = self.scale_table.device
        self.scale_table = torch.as_tensor(scale_table, dtype=torch.float32).to(dev)
        self.update()
        return True

    def update(self):
        st = self.scale_table.detach().cpu().numpy().astype(np.float32)
        cdf, length, offset = ops.gaussian_cdf_tables(st, self.tail_mass)
        self._set_tables(cdf, length, offset)

    def _table_dev(self, device):
        key = str(device)
        hit = self._dev_cache.get(key)
        if hit is None:
            with ops.cache_lock:
                hit = self.scale_table.detach().float().contiguous().to(device)
                ops.publish(hit)
                self._dev_cache[key] = hit
        return hit

    @staticmethod
    def _params(scales, means):
        s = _rows_cn(scales)
        m = _rows_cn(means) if means is not None else torch.zeros_like(s)
        return torch.cat([s, m], dim=1)

    def build_indexes(self, scales):
        s = _rows_cn(scales)
        params = torch.cat([s, s], dim=1)
        idx = ops.gc_build_indexes(params, None, s.shape[0], s.shape[1], self._table_dev(s.device), self._scale_bound)
        return idx.unsqueeze(0)

    def forward_rows(self, y, params, rows=None, training=None):
        """Row-major form of forward: y [N, C], params [P, 2C] = scales | means, rows: latent row ->
        params row.  -> (y_hat or y + noise [N, C], likelihood [N, C])."""
        from . import autograd as AG
        lb = self._likelihood_bound if self.use_likelihood_bound else 0.0
        if training if training is not None else self.training:
            y_noisy = y + quantization_noise(y)
            return y_noisy, AG.gc_likelihood(y_noisy, params, rows, self._scale_bound, lb)
        y_hat, lik = ops.gc_likelihood(y, params, rows, self._scale_bound, lb)
        return y_hat, lik.t()

    def forward(self, inputs, scales, means=None, training=None):
        if training if training is not None else self.training:
            y_noisy, lik = self.forward_rows(_rows_cn(inputs), self._params(scales, means), training=True)
            return y_noisy.t().unsqueeze(0), lik.t().unsqueeze(0)
        y = _rows_cn(inputs)
        y_hat, lik = ops.gc_likelihood(y, self._params(scales, means), None, self._scale_bound,
                                       self._likelihood_bound if self.use_likelihood_bound else 0.0)
        return y_hat.t().unsqueeze(0), lik.unsqueeze(0)

    def compress(self, inputs, indexes, means=None):
        y = _rows_cn(inputs)
        m = _rows_cn(means) if means is not None else torch.zeros_like(y)
        sym, _ = ops.gc_quantize_index(y, torch.cat([m, m], dim=1), None, self._table_dev(y.device), self._scale_bound)
        return [self._encode(sym, indexes[0].contiguous())]

    def compress_fused_gen(self, y, params, rows):
        """Fused path: y [N, C], params [Np, 2C] (scales | means), rows [N] -> (bytes, symbols, indexes).
        Generator: yields Wait objects (symbols reaching the host, the coder finishing)."""
        self._tables()  # materialise the host tables on the calling thread
        sym, idx = ops.gc_quantize_index(y, params, rows, self._table_dev(y.device), self._scale_bound)
        (sym_h, idx_h), ev = _d2h_async(sym, idx)
        yield Wait(event=ev)
        fut = _coder_pool.submit(self._encode_host, sym_h.numpy(), idx_h.numpy())
        yield Wait(future=fut)
        string = fut.result()
        _pinned_release(sym_h)
        _pinned_release(idx_h)
        return string, sym, idx

    def compress_fused(self, y, params, rows):
        return drive(self.compress_fused_gen(y, params, rows))

    def decompress(self, strings, indexes, dtype=torch.float, means=None):
        idx_dev = indexes[0].contiguous()
        sym = self._decode(strings[0], _to_host(idx_dev)).reshape(idx_dev.shape)
        sym_dev = _to_dev(sym, idx_dev.device)
        c, n = sym.shape
        m = _rows_cn(means) if means is not None else torch.zeros((n, c), device=idx_dev.device)
        y_hat = ops.gc_dequantize(sym_dev, torch.cat([m, m], dim=1), None)
        return y_hat.t().unsqueeze(0)

    def decompress_fused_gen(self, string, params, rows, n, c):
        """Fused path -> y_hat [N, C].  Generator (see compress_fused_gen)."""
        self._tables()
        idx = ops.gc_build_indexes(params, rows, n, c, self._table_dev(params.device), self._scale_bound)
        (idx_h,), ev = _d2h_async(idx)
        yield Wait(event=ev)
        fut = _coder_pool.submit(self._decode, string, idx_h.numpy())
        yield Wait(future=fut)
        sym = fut.result().reshape(c, n)
        _pinned_release(idx_h)
        return ops.gc_dequantize(_h2d_async(sym, params.device), params, rows)

    def decompress_fused(self, string, params, rows, n, c):
        return drive(self.decompress_fused_gen(string, params, rows, n, c))


class NonNegativeParametrizer(nn.Module):
    """compressai.ops.parametrizers.NonNegativeParametrizer: x -> max(x, bound)^2 - pedestal."""

    def __init__(self, minimum=0.0, reparam_offset=2 ** -18):
        super().__init__()
        pedestal = float(reparam_offset) ** 2
        self.register_buffer("pedestal", torch.Tensor([pedestal]))
        self.lower_bound = LowerBound((float(minimum) + pedestal) ** 0.5)

    def init(self, x):
        return torch.sqrt(torch.max(x + self.pedestal, self.pedestal))

    def forward(self, x):
        out = self.lower_bound(x)
        return out ** 2 - self.pedestal


class GDN(nn.Module):
    """compressai.layers.GDN (model/blocks.py:258-260 subclasses it as MinkowskiGDN; ScaledBlock constructs
    one, :27, and nothing on the codec path calls it): parameters ``beta`` [C], ``gamma`` [C, C] and the two
    parametrizers, with CompressAI's dense forward y = x / sqrt(beta + gamma * x^2) over [B, C, ...]."""

    def __init__(self, in_channels, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        self.inverse = bool(inverse)
        self.beta_reparam = NonNegativeParametrizer(minimum=float(beta_min))
        self.beta = nn.Parameter(self.beta_reparam.init(torch.ones(int(in_channels))))
        self.gamma_reparam = NonNegativeParametrizer()
        self.gamma = nn.Parameter(self.gamma_reparam.init(float(gamma_init) * torch.eye(int(in_channels))))

    def forward(self, x):
        C = x.shape[1]
        beta = self.beta_reparam(self.beta)
        gamma = self.gamma_reparam(self.gamma).reshape(C, C, *([1] * (x.dim() - 2)))
        conv = {3: torch.nn.functional.conv1d, 4: torch.nn.functional.conv2d, 5: torch.nn.functional.conv3d}[x.dim()]
        norm = conv(x ** 2, gamma, beta)
        norm = torch.sqrt(norm) if self.inverse else torch.rsqrt(norm)
        return x * norm


class CompressionModel(nn.Module):
    """compressai.models.base.CompressionModel: update(force=) walks the entropy sub-modules,
    aux_loss() sums the bottleneck losses (model/model.py:15,34,45)."""

    def update(self, scale_table=None, force=False):
        if scale_table is None:
            scale_table = get_scale_table()
        updated = False
        for _, m in self.named_modules():
            if isinstance(m, EntropyBottleneck):
                updated |= m.update(force=force)
            if isinstance(m, GaussianConditional):
                updated |= m.update_scale_table(scale_table, force=force)
        return updated

    def aux_loss(self):
        return sum(m.loss() for m in self.modules() if isinstance(m, EntropyBottleneck))
