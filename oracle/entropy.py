"""ORACLE — TEST INFRASTRUCTURE ONLY.  **parity unpinned**.

Entropy stage of the reference path, restating CompressAI 1.2.4 (requirements.txt:9, not
installable here) per SURVEY.md Appendix B [upstream-recalled]:
``EntropyBottleneck`` (model/entropy_models.py:269,313,352-353,393), ``GaussianConditional``
(model/entropy_models.py:270,330,371-372,407-408), the rANS coder and ``pmf_to_quantized_cdf``.

Tables are evaluated in float64 and rounded once to float32 before ``pmf_to_quantized_cdf``
(CompressAI evaluates them in float32 torch ops on whatever device the model sits on, so the
upstream tables are themselves device-dependent).
"""
import ctypes

import numpy as np
from scipy.special import erfc, expit
from scipy.stats import norm

from . import prims

_p = prims._p


# ------------------------------------------------------------------ coder
def pmf_to_quantized_cdf(pmf, precision=16):
    pmf = np.ascontiguousarray(pmf, dtype=np.float32)
    cdf = np.empty(pmf.shape[0] + 1, dtype=np.uint32)
    rc = prims.lib().orc_pmf_to_quantized_cdf(_p(pmf), pmf.shape[0], precision, _p(cdf))
    if rc != 0:
        raise ValueError(f"pmf_to_quantized_cdf failed ({rc})")
    return cdf.astype(np.int32)


def pmf_to_cdf_table(pmf, tail_mass, pmf_length, max_length):
    """EntropyModel._pmf_to_cdf [upstream-recalled]: one padded int32 row per table."""
    cdf = np.zeros((len(pmf_length), max_length + 2), dtype=np.int32)
    for i in range(len(pmf_length)):
        prob = np.concatenate([pmf[i, :pmf_length[i]], tail_mass[i].reshape(1)]).astype(np.float32)
        c = pmf_to_quantized_cdf(prob, 16)
        cdf[i, :c.shape[0]] = c
    return cdf


def rans_encode(symbols, indexes, cdf, cdf_length, offset):
    symbols = np.ascontiguousarray(symbols, dtype=np.int32).reshape(-1)
    indexes = np.ascontiguousarray(indexes, dtype=np.int32).reshape(-1)
    cdf = np.ascontiguousarray(cdf, dtype=np.int32)
    cdf_length = np.ascontiguousarray(cdf_length, dtype=np.int32)
    offset = np.ascontiguousarray(offset, dtype=np.int32)
    n = symbols.shape[0]
    cap = 8 * n + 4096
    while True:
        out = np.empty(cap, dtype=np.uint8)
        nb = prims.lib().orc_rans_encode(_p(symbols), _p(indexes), n, _p(cdf), cdf.shape[1],
                                         _p(cdf_length), _p(offset), cdf.shape[0], _p(out), cap)
        if nb == -1:
            cap *= 4
            continue
        if nb < 0:
            raise ValueError(f"rans_encode failed ({nb})")
        return out[:nb].tobytes()


def rans_decode(string, indexes, cdf, cdf_length, offset):
    indexes = np.ascontiguousarray(indexes, dtype=np.int32).reshape(-1)
    cdf = np.ascontiguousarray(cdf, dtype=np.int32)
    cdf_length = np.ascontiguousarray(cdf_length, dtype=np.int32)
    offset = np.ascontiguousarray(offset, dtype=np.int32)
    enc = np.frombuffer(string, dtype=np.uint8)
    out = np.empty(indexes.shape[0], dtype=np.int32)
    rc = prims.lib().orc_rans_decode(_p(enc), enc.shape[0], _p(indexes), indexes.shape[0], _p(cdf),
                                     cdf.shape[1], _p(cdf_length), _p(offset), cdf.shape[0], _p(out))
    if rc != 0:
        raise ValueError(f"rans_decode failed ({rc})")
    return out


def rans_encode_py(symbols, indexes, cdf, cdf_length, offset):
    """Pure-Python big-integer restatement of the same coder for tiny known-answer tests."""
    L, syms = 1 << 31, []
    for s, idx in zip(symbols, indexes):
        c, max_value = cdf[idx], int(cdf_length[idx]) - 2
        v, raw = int(s) - int(offset[idx]), 0
        if v < 0:
            raw, v = -2 * v - 1, max_value
        elif v >= max_value:
            raw, v = 2 * (v - max_value), max_value
        syms.append((int(c[v]), int(c[v + 1]) - int(c[v]), False))
        if v == max_value:
            nb = 0
            while (raw >> (nb * 4)) != 0:
                nb += 1
            val = nb
            while val >= 15:
                syms.append((15, 16, True))
                val -= 15
            syms.append((val, val + 1, True))
            for j in range(nb):
                val = (raw >> (4 * j)) & 15
                syms.append((val, val + 1, True))
    x, words = L, []
    for start, freq, bypass in reversed(syms):
        if not bypass:
            if x >= ((L >> 16) << 32) * freq:
                words.append(x & 0xFFFFFFFF)
                x >>= 32
            x = ((x // freq) << 16) + (x % freq) + start
        else:
            if x >= ((L >> 16) << 32) * (1 << 12):
                words.append(x & 0xFFFFFFFF)
                x >>= 32
            x = (x << 4) | start
    words.append(x >> 32)
    words.append(x & 0xFFFFFFFF)
    return np.array(words[::-1], dtype="<u4").tobytes()


# ------------------------------------------------------------------ GaussianConditional
SCALE_MIN, SCALE_MAX, SCALE_LEVELS = 0.11, 256.0, 64


def get_scale_table():
    """compressai.models.utils.get_scale_table, literally: ``torch.exp(torch.linspace(math.log(min),
    math.log(max), levels))`` — float32 linspace and float32 exp IN TORCH.  A float64 numpy evaluation rounded to
    float32 differs from it in the last bit of 46 of the 64 entries, and a scale that falls between the two
    versions of a threshold then gets a different CDF index: found by the full-frame parity test
    (tests/test_gpu_baseline_sizes.py: 1 index of 2.3 M on the 799,458-point ablation frame)."""
    import math
    import torch
    return torch.exp(torch.linspace(math.log(SCALE_MIN), math.log(SCALE_MAX), SCALE_LEVELS)).numpy().astype(np.float32)


class GaussianConditionalOracle:
    """GaussianConditional(None) after update_scale_table(get_scale_table()) (Appendix B2)."""

    scale_bound = np.float32(0.11)
    tail_mass = 1e-9
    likelihood_bound = np.float32(1e-9)

    def __init__(self):
        self.scale_table = get_scale_table()
        self.update()

    def update(self):
        t = self.scale_table.astype(np.float64)
        multiplier = -norm.ppf(self.tail_mass / 2)
        pmf_center = np.ceil(self.scale_table * np.float32(multiplier)).astype(np.int32)  # float32 product, as torch
        pmf_length = 2 * pmf_center + 1
        max_length = int(pmf_length.max())
        samples = np.abs(np.arange(max_length)[None, :] - pmf_center[:, None]).astype(np.float64)
        cum = lambda v: 0.5 * erfc(-(2 ** -0.5) * v)
        upper = cum((0.5 - samples) / t[:, None])
        lower = cum((-0.5 - samples) / t[:, None])
        pmf = (upper - lower).astype(np.float32)
        tail = (2 * lower[:, :1]).astype(np.float32)
        self.quantized_cdf = pmf_to_cdf_table(pmf, tail, pmf_length, max_length)
        self.offset = (-pmf_center).astype(np.int32)
        self.cdf_length = (pmf_length + 2).astype(np.int32)

    def build_indexes(self, scales):
        s = np.maximum(np.asarray(scales, dtype=np.float32), self.scale_bound)
        idx = np.full(s.shape, SCALE_LEVELS - 1, dtype=np.int32)
        for t in self.scale_table[:-1]:
            idx -= (s <= t).astype(np.int32)
        return idx

    @staticmethod
    def quantize(x, means):
        """EntropyModel.quantize(mode='symbols'): torch.round(x - means).int() (half to even)."""
        return np.rint(np.asarray(x, np.float32) - np.asarray(means, np.float32)).astype(np.int32)

    def compress(self, x, indexes, means):
        """x, indexes, means: [C, N] (the [1,C,N] batch element); channel-major flattening."""
        return rans_encode(self.quantize(x, means), indexes, self.quantized_cdf, self.cdf_length,
                           self.offset)

    def decompress(self, string, indexes, means):
        sym = rans_decode(string, indexes, self.quantized_cdf, self.cdf_length, self.offset)
        return sym.reshape(indexes.shape).astype(np.float32) + np.asarray(means, np.float32)

    def forward(self, x, scales, means):
        """Eval-mode __call__: (dequantised y_hat, likelihood) — fp64 evaluation, tolerance oracle."""
        x, means = np.asarray(x, np.float32), np.asarray(means, np.float32)
        y_hat = np.rint(x - means).astype(np.float32) + means
        s = np.maximum(np.asarray(scales, np.float32), self.scale_bound).astype(np.float64)
        v = np.abs((y_hat - means).astype(np.float64))
        cum = lambda u: 0.5 * erfc(-(2 ** -0.5) * u)
        lik = cum((0.5 - v) / s) - cum((-0.5 - v) / s)
        return y_hat, np.maximum(lik, float(self.likelihood_bound)).astype(np.float32)


# ------------------------------------------------------------------ EntropyBottleneck
def _softplus(v):
    return np.logaddexp(0.0, v)


class EntropyBottleneckOracle:
    """EntropyBottleneck(channels), filters (3,3,3,3), init_scale 10, tail_mass 1e-9 (Appendix B1).
    params: dict with _matrix{i} [C,f_{i+1},f_i], _bias{i} [C,f_{i+1},1], _factor{i} (i<4), quantiles [C,1,3]."""

    likelihood_bound = np.float32(1e-9)

    def __init__(self, params):
        self.p = {k: np.asarray(v, dtype=np.float32) for k, v in params.items()}
        self.C = self.p["quantiles"].shape[0]
        self.update()

    def logits_cumulative(self, v):
        """v: [C, 1, M] float64."""
        logits = v
        for i in range(5):
            logits = np.matmul(_softplus(self.p[f"_matrix{i}"].astype(np.float64)), logits)
            logits = logits + self.p[f"_bias{i}"].astype(np.float64)
            if i < 4:
                logits = logits + np.tanh(self.p[f"_factor{i}"].astype(np.float64)) * np.tanh(logits)
        return logits

    def medians(self):
        return self.p["quantiles"][:, 0, 1].copy()

    def update(self):
        q = self.p["quantiles"]
        med = q[:, 0, 1]
        minima = np.maximum(np.ceil(med - q[:, 0, 0]).astype(np.int32), 0)
        maxima = np.maximum(np.ceil(q[:, 0, 2] - med).astype(np.int32), 0)
        self.offset = (-minima).astype(np.int32)
        pmf_start = med - minima.astype(np.float32)
        pmf_length = maxima + minima + 1
        max_length = int(pmf_length.max())
        samples = (np.arange(max_length, dtype=np.float32)[None, None, :] + pmf_start[:, None, None])
        samples = samples.astype(np.float64)
        lower = self.logits_cumulative(samples - 0.5)
        upper = self.logits_cumulative(samples + 0.5)
        sign = -np.sign(lower + upper)
        pmf = np.abs(expit(sign * upper) - expit(sign * lower))[:, 0, :].astype(np.float32)
        tail = (expit(lower[:, 0, :1]) + expit(-upper[:, 0, -1:])).astype(np.float32)
        self.quantized_cdf = pmf_to_cdf_table(pmf, tail, pmf_length, max_length)
        self.cdf_length = (pmf_length + 2).astype(np.int32)

    def _indexes(self, n):
        return np.repeat(np.arange(self.C, dtype=np.int32)[:, None], n, axis=1)

    def compress(self, x):
        """x: [C, N]."""
        sym = np.rint(np.asarray(x, np.float32) - self.medians()[:, None]).astype(np.int32)
        return rans_encode(sym, self._indexes(x.shape[1]), self.quantized_cdf, self.cdf_length,
                           self.offset)

    def decompress(self, string, n):
        sym = rans_decode(string, self._indexes(n), self.quantized_cdf, self.cdf_length, self.offset)
        return sym.reshape(self.C, n).astype(np.float32) + self.medians()[:, None]

    def forward(self, x):
        """Eval-mode __call__: (z_hat, likelihood)."""
        med = self.medians()[:, None]
        z_hat = np.rint(np.asarray(x, np.float32) - med).astype(np.float32) + med
        v = z_hat.astype(np.float64)[:, None, :]
        lower = self.logits_cumulative(v - 0.5)
        upper = self.logits_cumulative(v + 0.5)
        sign = -np.sign(lower + upper)
        lik = np.abs(expit(sign * upper) - expit(sign * lower))[:, 0, :]
        return z_hat, np.maximum(lik, float(self.likelihood_bound)).astype(np.float32)
