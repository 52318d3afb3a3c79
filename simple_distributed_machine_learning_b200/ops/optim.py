"""Per-stage fused optimizers (``csrc/optim.cu``).

The reference builds a ``DistributedOptimizer(optim.SGD, lr=0.1, momentum=0.5)`` over
RRefs to the parameters of both ranks and fans out one RPC per owner per step
(/root/reference/simple_distributed.py:100-104, :113).  Here every rank owns a local
optimizer over its own stage's parameters -- the same natural partitioning, without RPC
-- and the update of all tensors is ONE kernel launch that also un-scales the gradient,
refreshes the bf16 compute copy of the weights and clears the gradient accumulator.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Iterable, List, Optional

import torch

from . import _lib


class _FusedBase:
    def __init__(self, params: Iterable[torch.Tensor], grads: Optional[List[torch.Tensor]] = None,
                 lowp: Optional[List[Optional[torch.Tensor]]] = None):
        self.params: List[torch.Tensor] = list(params)
        for p in self.params:
            assert p.dtype == torch.float32 and p.is_contiguous(), "fp32 contiguous master params"
        # explicit gradient buffers (manual stages) or .grad (autograd stages, resolved lazily)
        self.grads = grads
        self.lowp = lowp
        self.step_count = 0
        self._tables = None

    def _grad_list(self) -> List[torch.Tensor]:
        if self.grads is not None:
            return self.grads
        out = []
        for p in self.params:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            out.append(p.grad)
        return out

    @staticmethod
    def _ptr_array(tensors: List[Optional[torch.Tensor]]):
        arr = (ctypes.c_void_p * len(tensors))()
        for i, t in enumerate(tensors):
            arr[i] = t.data_ptr() if t is not None else None
        return arr

    def zero_grad(self) -> None:
        if self.grads is not None:
            for g in self.grads:
                g.zero_()
        else:
            for p in self.params:
                if p.grad is not None:
                    p.grad.zero_()


class FusedSGD(_FusedBase):
    """``torch.optim.SGD`` semantics (momentum, no dampening / nesterov), one launch."""

    def __init__(self, params, lr: float, momentum: float = 0.0, weight_decay: float = 0.0,
                 grads=None, lowp=None):
        super().__init__(params, grads, lowp)
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.momentum_buffers: List[torch.Tensor] = [torch.zeros_like(p) for p in self.params]

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0, zero_grad: bool = False) -> None:
        grads = self._grad_list()
        # momentum buffers start at zero, so "buf = g" on the first step is the same as the
        # steady-state "buf = momentum * buf + g": no step-dependent branch, which keeps the
        # launch identical from step to step (CUDA-graph friendly).
        first = False
        if self.params and self.params[0].is_cuda:
            n = len(self.params)
            sizes = (ctypes.c_longlong * n)(*[p.numel() for p in self.params])
            rc = _lib.cuda_lib().sdml_sgd_multi(
                n, self._ptr_array(self.params), self._ptr_array(grads),
                self._ptr_array(self.momentum_buffers),
                self._ptr_array(self.lowp) if self.lowp is not None else None, sizes,
                float(self.lr), float(self.momentum), float(self.weight_decay), float(grad_scale),
                int(first), int(zero_grad), torch.cuda.current_stream().cuda_stream)
            _lib.check(rc, "sdml_sgd_multi")
        else:
            for i, (p, g) in enumerate(zip(self.params, grads)):
                d = g * grad_scale + self.weight_decay * p
                if self.momentum != 0.0:
                    buf = self.momentum_buffers[i]
                    if first:
                        buf.copy_(d)
                    else:
                        buf.mul_(self.momentum).add_(d)
                    d = buf
                p.add_(d, alpha=-self.lr)
                if zero_grad:
                    g.zero_()
                if self.lowp is not None and self.lowp[i] is not None:
                    self.lowp[i].copy_(p)
        self.step_count += 1

    def state_dict(self) -> Dict:
        return {"kind": "sgd", "step": self.step_count, "lr": self.lr, "momentum": self.momentum,
                "weight_decay": self.weight_decay,
                "momentum_buffers": [b.detach().cpu().clone() for b in self.momentum_buffers]}

    def load_state_dict(self, sd: Dict) -> None:
        assert sd["kind"] == "sgd"
        self.step_count = int(sd["step"])
        for b, s in zip(self.momentum_buffers, sd["momentum_buffers"]):
            b.copy_(s)


class FusedAdam(_FusedBase):
    """``torch.optim.Adam`` semantics (L2 weight decay, bias correction), one launch."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, grads=None, lowp=None):
        super().__init__(params, grads, lowp)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0, zero_grad: bool = False) -> None:
        grads = self._grad_list()
        self.step_count += 1
        b1, b2 = self.betas
        if self.params and self.params[0].is_cuda:
            n = len(self.params)
            sizes = (ctypes.c_longlong * n)(*[p.numel() for p in self.params])
            rc = _lib.cuda_lib().sdml_adam_multi(
                n, self._ptr_array(self.params), self._ptr_array(grads),
                self._ptr_array(self.exp_avg), self._ptr_array(self.exp_avg_sq),
                self._ptr_array(self.lowp) if self.lowp is not None else None, sizes,
                float(self.lr), float(b1), float(b2), float(self.eps), float(self.weight_decay),
                float(grad_scale), int(self.step_count), int(zero_grad),
                torch.cuda.current_stream().cuda_stream)
            _lib.check(rc, "sdml_adam_multi")
        else:
            bias1 = 1.0 - b1 ** self.step_count
            bias2 = 1.0 - b2 ** self.step_count
            for i, (p, g) in enumerate(zip(self.params, grads)):
                d = g * grad_scale + self.weight_decay * p
                self.exp_avg[i].mul_(b1).add_(d, alpha=1.0 - b1)
                self.exp_avg_sq[i].mul_(b2).addcmul_(d, d, value=1.0 - b2)
                denom = self.exp_avg_sq[i].sqrt() / (bias2 ** 0.5) + self.eps
                p.addcdiv_(self.exp_avg[i], denom, value=-self.lr / bias1)
                if zero_grad:
                    g.zero_()
                if self.lowp is not None and self.lowp[i] is not None:
                    self.lowp[i].copy_(p)

    def state_dict(self) -> Dict:
        return {"kind": "adam", "step": self.step_count, "lr": self.lr, "betas": self.betas,
                "eps": self.eps, "weight_decay": self.weight_decay,
                "exp_avg": [t.detach().cpu().clone() for t in self.exp_avg],
                "exp_avg_sq": [t.detach().cpu().clone() for t in self.exp_avg_sq]}

    def load_state_dict(self, sd: Dict) -> None:
        assert sd["kind"] == "adam"
        self.step_count = int(sd["step"])
        for t, s in zip(self.exp_avg, sd["exp_avg"]):
            t.copy_(s)
        for t, s in zip(self.exp_avg_sq, sd["exp_avg_sq"]):
            t.copy_(s)


def make_optimizer(kind: str, params, lr: float, momentum: float = 0.0, weight_decay: float = 0.0,
                   grads=None, lowp=None):
    if kind == "sgd":
        return FusedSGD(params, lr=lr, momentum=momentum, weight_decay=weight_decay, grads=grads, lowp=lowp)
    if kind == "adam":
        return FusedAdam(params, lr=lr, weight_decay=weight_decay, grads=grads, lowp=lowp)
    raise ValueError(f"unknown optimizer {kind!r}")
