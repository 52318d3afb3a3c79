"""Fused softmax cross-entropy (``csrc/softmax_ce.cu``) front-end.

Replaces the reference's split ``F.log_softmax`` on rank 1
(/root/reference/simple_distributed.py:79) + ``F.nll_loss`` on rank 0 (ref:111, ref:126)
+ the eval arg-max/eq/sum (ref:127-128) by one kernel on the last stage that returns the
loss sum, the logits gradient and the number of correct predictions.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib


def softmax_ce(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float = 1.0,
               want_grad: bool = True, want_correct: bool = False,
               loss_sum: Optional[torch.Tensor] = None, correct: Optional[torch.Tensor] = None,
               dlogits: Optional[torch.Tensor] = None,
               ) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor]]:
    """Returns ``(loss_sum[1] fp32, dlogits or None, correct[1] int32 or None)``.

    ``loss_sum`` / ``correct`` are *accumulated into* when passed in (so a whole epoch of
    eval batches needs one D2H read); ``dlogits = (softmax - onehot) * grad_scale``.
    """
    assert logits.dim() == 2 and targets.dim() == 1 and targets.shape[0] == logits.shape[0]
    rows, classes = logits.shape
    if loss_sum is None:
        loss_sum = torch.zeros(1, dtype=torch.float32, device=logits.device)
    if want_correct and correct is None:
        correct = torch.zeros(1, dtype=torch.int32, device=logits.device)
    if logits.is_cuda:
        assert logits.stride(1) == 1 and targets.dtype == torch.int64 and targets.is_contiguous()
        dtype_code = {torch.float32: 0, torch.bfloat16: 1}[logits.dtype]
        if want_grad and dlogits is None:
            dlogits = torch.empty_like(logits)
        rc = _lib.cuda_lib().sdml_softmax_ce(
            logits.data_ptr(), logits.stride(0), dtype_code, targets.data_ptr(), rows, classes,
            loss_sum.data_ptr(), dlogits.data_ptr() if want_grad else None,
            dlogits.stride(0) if want_grad else 0, float(grad_scale),
            correct.data_ptr() if want_correct else None, torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_softmax_ce")
        return loss_sum, (dlogits if want_grad else None), (correct if want_correct else None)
    # CPU oracle path (gloo plumbing configuration, unit tests)
    x = logits.float()
    lse = torch.logsumexp(x, dim=1)
    loss_sum += (lse - x.gather(1, targets[:, None]).squeeze(1)).sum()
    if want_correct:
        correct += (x.argmax(dim=1) == targets).sum().to(torch.int32)
    if want_grad:
        p = torch.softmax(x, dim=1)
        p[torch.arange(rows), targets] -= 1.0
        g = (p * grad_scale).to(logits.dtype)
        if dlogits is not None:
            dlogits.copy_(g)
        else:
            dlogits = g
    return loss_sum, (dlogits if want_grad else None), (correct if want_correct else None)


class _SoftmaxCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, scale):
        loss_sum, dlogits, _ = softmax_ce(logits.detach(), targets, grad_scale=scale)
        ctx.save_for_backward(dlogits)
        return (loss_sum * scale).squeeze(0)

    @staticmethod
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits * grad_out, None, None


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, scale: float) -> torch.Tensor:
    """Autograd-visible ``scale * sum_r CE_r`` (``scale = 1/batch`` gives the mean loss of
    ref:111) whose backward is the gradient already produced by the fused forward."""
    return _SoftmaxCEFn.apply(logits, targets, scale)
