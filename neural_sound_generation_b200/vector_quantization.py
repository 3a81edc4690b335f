"""Drop-in for /root/reference/src/vector_quantization.py: ``vq`` and ``vq_st``.

Same names, call signatures, return shapes / dtypes and autograd behaviour as
the reference's two ``torch.autograd.Function``s; the arithmetic runs in the
sm_100a kernels behind ``nsg_vq_forward`` / ``nsg_scatter_add``.  CUDA tensors
only -- a CPU tensor raises (there is no fallback path).
"""
import torch
from torch.autograd import Function

from . import ops

# None = automatic: the tcgen05 split-bf16 tensor-core search (mode 1: top-2 candidates re-scored exactly in fp32,
# lowest index on ties) whenever it serves the shape, else the exact fp32 CUDA-core search (mode 0).  Set to 0 / 1 to pin.
DEFAULT_MODE = None


def _mode(codebook):
    if DEFAULT_MODE is not None:
        return DEFAULT_MODE
    return 1 if ops.vq_mode_serves(codebook.size(0), codebook.size(1), 1) else 0


def _prep(inputs, codebook):
    if not inputs.is_cuda or not codebook.is_cuda:
        raise ops.A.NsgError('vq/vq_st: CUDA tensors required (no CPU path in neural_sound_generation_b200)')
    if inputs.dtype not in (torch.float32, torch.bfloat16):
        raise ops.A.NsgError('vq/vq_st: inputs must be float32 or bfloat16, got %s' % inputs.dtype)
    if not inputs.is_contiguous():
        # the reference calls inputs.view(-1, D) (vector_quantization.py:10), which raises on this too
        raise RuntimeError('vq: inputs must be contiguous (the reference uses .view(-1, D))')
    cb = codebook.detach()
    if cb.dtype != torch.float32:
        cb = cb.float()
    return cb.contiguous()


class VectorQuantization(Function):
    """indices = argmin_k |inputs - codebook[k]|^2 (vector_quantization.py:4-30)."""

    @staticmethod
    def forward(ctx, inputs, codebook):
        cb = _prep(inputs, codebook)
        d = cb.size(1)
        if inputs.size(-1) != d:
            raise RuntimeError('vq: last dim of inputs (%d) != embedding size (%d)' % (inputs.size(-1), d))
        n = inputs.numel() // d
        indices_flatten = torch.empty((n,), dtype=torch.int64, device=inputs.device)
        ops.vq_forward(inputs.detach(), cb, indices_flatten, None, None, mode=_mode(cb))
        indices = indices_flatten.view(*inputs.size()[:-1])
        ctx.mark_non_differentiable(indices)
        return indices

    @staticmethod
    def backward(ctx, grad_output):
        raise RuntimeError('Trying to call `.grad()` on graph containing '
                           '`VectorQuantization`. The function `VectorQuantization` '
                           'is not differentiable. Use `VectorQuantizationStraightThrough` '
                           'if you want a straight-through estimator of the gradient.')


class VectorQuantizationStraightThrough(Function):
    """(codes, indices_flatten) with straight-through gradient (vector_quantization.py:32-63)."""

    @staticmethod
    def forward(ctx, inputs, codebook):
        cb = _prep(inputs, codebook)
        d = cb.size(1)
        n = inputs.numel() // d
        indices_flatten = torch.empty((n,), dtype=torch.int64, device=inputs.device)
        codes = torch.empty_like(inputs)
        ops.vq_forward(inputs.detach(), cb, indices_flatten, codes, None, mode=_mode(cb))
        ctx.save_for_backward(indices_flatten, codebook)
        ctx.mark_non_differentiable(indices_flatten)
        return (codes, indices_flatten)

    @staticmethod
    def backward(ctx, grad_output, grad_indices):
        grad_inputs, grad_codebook = None, None
        if ctx.needs_input_grad[0]:
            grad_inputs = grad_output.clone()                     # straight-through (:52)
        if ctx.needs_input_grad[1]:
            indices, codebook = ctx.saved_tensors
            g = grad_output.contiguous()                          # (:58)
            table = torch.zeros(codebook.shape, dtype=torch.float32, device=codebook.device)
            ops.scatter_add(g, indices, table, alpha=1.0)         # index_add_ (:61)
            grad_codebook = table if codebook.dtype == torch.float32 else table.to(codebook.dtype)
        return (grad_inputs, grad_codebook)


vq = VectorQuantization.apply
vq_st = VectorQuantizationStraightThrough.apply
__all__ = ['vq', 'vq_st', 'VectorQuantization', 'VectorQuantizationStraightThrough']
