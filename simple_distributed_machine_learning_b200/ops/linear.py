"""Linear (+bias, +ReLU, +dropout) autograd front-ends.

* :func:`linear_f32` -- exact-fp32 small-shape path (``csrc/linear_f32.cu``) used by the
  reference model's fc1/fc2 (/root/reference/simple_distributed.py:63-64, 75-77).
* The bf16 tensor-core path is driven explicitly (no autograd) by
  :class:`..models.mlp.MLPStage` through :func:`..ops.gemm.gemm_bf16`, because the
  zero-bubble schedules need dgrad and wgrad as separately schedulable ops.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _lib


class _LinearF32Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, relu, drop_p, seed, offset, offset_dev):
        x = x.contiguous()
        M, K = x.shape
        N = w.shape[0]
        y = torch.empty(M, N, dtype=torch.float32, device=x.device)
        rc = _lib.cuda_lib().sdml_linear_f32_fwd(
            x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, y.data_ptr(),
            M, N, K, int(relu), float(drop_p), int(seed), int(offset),
            offset_dev.data_ptr() if offset_dev is not None else None,
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_linear_f32_fwd")
        ctx.save_for_backward(x, w, y)
        ctx.offset_dev = offset_dev
        ctx.cfg = (bool(relu), float(drop_p), int(seed), int(offset), b is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        relu, drop_p, seed, offset, has_bias = ctx.cfg
        dy = dy.contiguous()
        M, K = x.shape
        N = w.shape[0]
        need_dx = ctx.needs_input_grad[0]
        dx = torch.empty_like(x) if need_dx else None
        dw = torch.empty_like(w)
        db = torch.empty(N, dtype=torch.float32, device=x.device) if has_bias else None
        rc = _lib.cuda_lib().sdml_linear_f32_bwd(
            x.data_ptr(), w.data_ptr(), y.data_ptr(), dy.data_ptr(),
            dx.data_ptr() if need_dx else None, dw.data_ptr(),
            db.data_ptr() if has_bias else None, M, N, K, int(relu), drop_p, seed, offset,
            ctx.offset_dev.data_ptr() if ctx.offset_dev is not None else None, 0,
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_linear_f32_bwd")
        return dx, dw, db, None, None, None, None, None


def linear_f32(x: torch.Tensor, w: torch.Tensor, b, relu: bool = False, drop_p: float = 0.0,
               seed: int = 0, offset: int = 0, offset_dev: torch.Tensor = None) -> torch.Tensor:
    """``dropout_p(relu?(x @ w.T + b))`` in fp32.  CUDA: hand-written kernels with a
    counter-based mask (regenerated in backward).  CPU: the plain PyTorch oracle.

    ``offset_dev`` (optional int64[1] device tensor) is added to ``offset`` on the device, so
    a CUDA-graph replay sees a fresh mask; it must not change between forward and backward.
    """
    if x.is_cuda:
        assert x.dtype == torch.float32 and w.dtype == torch.float32
        return _LinearF32Fn.apply(x, w, b, relu, drop_p, seed, offset, offset_dev)
    y = F.linear(x, w, b)
    if relu:
        y = F.relu(y)
    if drop_p > 0.0:
        y = F.dropout(y, p=drop_p, training=True)
    return y
