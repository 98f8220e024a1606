"""Differentiable forms of the path's operators (ColorModel.forward under
autograd, model/model.py:51-93, driven by train.py:189-203).

The reference gets its gradients from MinkowskiEngine's ConvolutionBackward /
ConvolutionTransposeBackward, MinkowskiPruning's backward, the interpolation backward behind
``features_at_coordinates`` and torch autograd over CompressAI's entropy modules.  Here:

  * ``sparse_conv``  — forward ``pcc_conv_fwd``; data gradient = ``pcc_conv_fwd`` over the REVERSED
    kernel map (built from the output side) with the per-offset transposed weights; weight gradient
    = ``pcc_conv_wgrad`` (deterministic); bias gradient = column sum.
  * ``gather``       — forward ``pcc_gather_rows``; backward ``pcc_scatter_rows_add``.
  * ``gc_likelihood``— ``pcc_gc_likelihood_train_fwd`` / ``_bwd`` (LowerBound gradient rule inside).

torch.autograd is the tape only; activation / FiLM / residual are plain tensor expressions in this
mode (the fused epilogue of the codec path has no backward).
"""
import torch

from . import ops
from ._lib import PccError


class _SparseConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, weight, bias, in_map, out_map, ksize, transpose):
        kmap = in_map.kernel_map(out_map, ksize, transpose=transpose)
        # precision policy of DESIGN.md §5: in the tensor-core mode the forward pass and the data gradient of
        # layers with cin, cout % 32 == 0 run on tcgen05 (tf32 operands, fp32 accumulation), and so does their
        # weight gradient; everything else on the fp32 CUDA-core kernels
        w3 = weight if weight.dim() == 3 else weight.unsqueeze(0)
        mode = ops.default_mode()
        if mode == ops.MODE_TF32 and not (ops.tf32_eligible(w3.shape[1], w3.shape[2]) and kmap.n_groups == 1):
            mode = ops.MODE_EXACT
        out = ops.conv(feats, kmap, weight, bias, mode=mode, round_out=False)
        ctx.save_for_backward(feats, weight)
        ctx.mode = mode
        ctx.maps = (in_map, out_map, ksize, transpose, kmap)
        ctx.has_bias = bias is not None
        ctx.bias_shape = None if bias is None else tuple(bias.shape)
        return out

    @staticmethod
    def backward(ctx, g):
        feats, weight = ctx.saved_tensors
        in_map, out_map, ksize, transpose, kmap = ctx.maps
        g = g.contiguous()
        w3 = weight if weight.dim() == 3 else weight.unsqueeze(0)
        K, cin, cout = w3.shape
        g_in = g_w = g_b = None
        if ctx.needs_input_grad[0]:
            # nbr[k][o] = i  <=>  rnbr[k][i] = o: the same offsets probed from the other side.  On ONE map with an
            # odd kernel the other side is the mirrored offset of the same kernel map (c_o = c_i - d_k = c_i +
            # d_{K-1-k}), so the data gradient runs over the forward map with the weights flipped along K and no
            # second map is built
            mirrored = in_map is out_map and ksize % 2 == 1 and not transpose
            rmap = kmap if mirrored else out_map.kernel_map(in_map, ksize, transpose=not transpose)
            mode = ctx.mode if (ctx.mode == ops.MODE_EXACT or rmap.n_groups == 1) else ops.MODE_EXACT
            if mode == ops.MODE_TF32:
                # the forward pass left the tf32-rounded [K, cout, cin] copy in the cache; its own tensor-core
                # layout is the rounded weight again (one small kernel, ordered by an event: no host wait)
                wt = ops.transposed_weight(weight)
                wt = wt.flip(0) if mirrored else wt
                wtt = ops.transposed_weight(wt)
            else:
                wt, wtt = w3.detach().transpose(1, 2), None
                wt = (wt.flip(0) if mirrored else wt).contiguous()
            g_in = ops.conv(g, rmap, wt, None, mode=mode, round_out=False, weight_t=wtt)
        if ctx.needs_input_grad[1]:
            g_w = ops.conv_wgrad(feats, g, kmap, cin, cout, mode=ctx.mode).view(weight.shape)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            g_b = (ops.col_sum(g) if g.shape[1] <= 1024 else g.sum(dim=0)).view(ctx.bias_shape)
        return g_in, g_w, g_b, None, None, None, None


class _Gather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, rows):
        ctx.rows, ctx.n_src = rows, feats.shape[0]
        return ops.gather_rows(feats, rows)

    @staticmethod
    def backward(ctx, g):
        return ops.scatter_rows_add(g, ctx.rows, ctx.n_src), None


class _GcLikelihood(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y_noisy, params, rows, scale_bound, lik_bound):
        ctx.save_for_backward(y_noisy, params)
        ctx.cfg = (rows, float(scale_bound), float(lik_bound))
        return ops.gc_likelihood_train_fwd(y_noisy, params, rows, scale_bound, lik_bound)

    @staticmethod
    def backward(ctx, g):
        y_noisy, params = ctx.saved_tensors
        rows, sb, lb = ctx.cfg
        g_y, g_prow = ops.gc_likelihood_train_bwd(y_noisy, params, rows, sb, lb, g)
        g_params = ops.scatter_rows_add(g_prow, rows, params.shape[0]) if ctx.needs_input_grad[1] else None
        return g_y, g_params, None, None, None


def wants_grad(*tensors):
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


def gather(feats, rows, add=None):
    """features_at_coordinates on-grid (gather-or-zero), differentiable in feats (and add)."""
    out = _Gather.apply(feats, rows) if rows is not None else feats
    return out if add is None else add + out


def gc_likelihood(y_noisy, params, rows, scale_bound, lik_bound):
    return _GcLikelihood.apply(y_noisy, params, rows, scale_bound, lik_bound)


def sparse_conv(feats, in_map, out_map, ksize, weight, bias=None, transpose=False, act=ops.ACT_NONE,
                film=None, film_rows=None, residual=None, res_rows=None):
    """The convolution with pcc_conv_fwd's epilogue order (bias, activation, FiLM, residual) written
    out as differentiable tensor expressions."""
    if not hasattr(in_map, "kernel_map") or type(in_map).__name__ == "SlabMap":
        raise PccError("the training path runs on unpartitioned maps (one process per batch shard)")
    v = _SparseConv.apply(feats, weight, bias, in_map, out_map, ksize, transpose)
    if act == ops.ACT_RELU:
        v = torch.relu(v)
    elif act == ops.ACT_LEAKY:
        v = torch.nn.functional.leaky_relu(v, ops.LEAKY_SLOPE)
    if film is not None:
        f = gather(film, film_rows)
        cf = f.shape[1] // 2
        v = v * f[:, :cf] + f[:, cf:]
    if residual is not None:
        v = v + gather(residual, res_rows)
    return v
