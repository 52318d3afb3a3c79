"""Stage runtimes: what one pipeline rank does for F / B / W ops.

The reference couples the stages through RRefs and distributed autograd
(/root/reference/simple_distributed.py:47-50, 71, 109-112).  Here each rank owns a
``Stage`` object with explicit per-micro-batch entry points that the schedule executor
(:mod:`.pipeline`) calls in schedule order:

* ``forward(mb, x, targets)``      -> activation for the next stage (``None`` on the last)
* ``backward_input(mb, dy)``       -> gradient for the previous stage (``None`` on the first)
* ``backward_weight(mb)``          -> accumulates weight gradients (no communication)
* ``end_step()``                   -> DP gradient reduction (if any) + optimizer step

Loss lives on the LAST stage (the reference computes it on rank 0, ref:111); the mean
over the global batch is obtained by scaling every micro-batch's loss by
``1 / global_batch`` so the accumulated gradient equals the reference's full-batch step.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from ..ops import loss as loss_ops
from ..ops.optim import make_optimizer


class Stage:
    """Interface + shared bookkeeping."""

    is_first: bool
    is_last: bool
    device: torch.device
    act_dtype: torch.dtype = torch.float32

    def __init__(self, is_first: bool, is_last: bool, device: torch.device):
        self.is_first, self.is_last, self.device = is_first, is_last, device
        self.loss_sum = torch.zeros(1, dtype=torch.float32, device=device)
        self.correct = torch.zeros(1, dtype=torch.int32, device=device)
        self._scale = 1.0
        self.dp_group = None
        self.dp_size = 1

    # ---- shapes of what crosses the stage boundary ----
    def in_shape(self, mb_size: int) -> Tuple[int, ...]:
        raise NotImplementedError

    def out_shape(self, mb_size: int) -> Tuple[int, ...]:
        raise NotImplementedError

    # ---- step protocol ----
    def begin_step(self, global_batch: int, sizes: Optional[List[int]] = None) -> None:
        """Called before the first op of a training step; ``global_batch`` = samples summed
        over all micro-batches (of this DP replica), ``sizes`` = rows of every micro-batch."""
        self._scale = 1.0 / float(global_batch)
        self.loss_sum.zero_()

    def input_buffer(self, mb: int) -> Optional[torch.Tensor]:
        """Optional destination for the received activation of ``mb`` (None: executor allocates)."""
        return None

    def grad_buffer(self, mb: int) -> Optional[torch.Tensor]:
        """Optional destination for the received gradient of ``mb`` (None: executor allocates)."""
        return None

    def forward(self, mb: int, x: Optional[torch.Tensor], targets: Optional[torch.Tensor]):
        raise NotImplementedError

    def backward_input(self, mb: int, dy: Optional[torch.Tensor]):
        raise NotImplementedError

    def backward_weight(self, mb: int) -> None:
        raise NotImplementedError

    def end_step(self) -> None:
        raise NotImplementedError

    def join_side_streams(self) -> None:
        """Make the current stream wait for work this stage put on its own side streams.  Called by
        executors that capture every op into its own CUDA graph (a capture must end with all forked
        streams joined); stages without side streams have nothing to do."""

    def mean_loss(self) -> torch.Tensor:
        """Mean training loss of the step (last stage only); device tensor, no sync."""
        return self.loss_sum * self._scale

    # ---- evaluation (no grad): accumulates loss_sum / correct ----
    def eval_begin(self) -> None:
        self.loss_sum.zero_()
        self.correct.zero_()

    def eval_forward(self, x: Optional[torch.Tensor], targets: Optional[torch.Tensor]):
        raise NotImplementedError

    # ---- checkpoint ----
    def state_dict(self) -> Dict[str, torch.Tensor]:
        raise NotImplementedError

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        raise NotImplementedError

    def set_training(self, training: bool) -> None:
        pass

    # ---- per-stage checkpoint files: {global stage index: state_dict} ----
    def checkpoint_parts(self, pp_idx: int) -> Dict[int, Dict[str, torch.Tensor]]:
        return {pp_idx: self.state_dict()}

    def load_checkpoint_parts(self, pp_idx: int, parts: Dict[int, Dict[str, torch.Tensor]]) -> None:
        self.load_state_dict(parts[pp_idx])

    def checkpoint_part_indices(self, pp_idx: int) -> List[int]:
        return [pp_idx]


class AutogradStage(Stage):
    """A stage whose math is an ``nn.Module`` differentiated by local autograd.

    B and W are not separable here: ``backward_input`` runs the whole local backward
    (parameter gradients accumulate in ``.grad``) and ``backward_weight`` is a no-op.
    Used for the reference LeNet stages.
    """

    def __init__(self, module: nn.Module, is_first: bool, is_last: bool, device: torch.device,
                 in_features: Tuple[int, ...], out_features: Tuple[int, ...], optimizer: str = "sgd",
                 lr: float = 0.1, momentum: float = 0.5, weight_decay: float = 0.0,
                 part_prefixes: Optional[Dict[int, str]] = None):
        super().__init__(is_first, is_last, device)
        self.module = module.to(device)
        # when several reference stages are co-located on one rank (pp=1), each still gets
        # its own checkpoint file: {reference stage index: sub-module prefix}
        self.part_prefixes = part_prefixes
        self._in, self._out = tuple(in_features), tuple(out_features)
        self.params = [p for p in self.module.parameters()]
        self.optimizer = make_optimizer(optimizer, [p.data for p in self.params], lr=lr,
                                        momentum=momentum, weight_decay=weight_decay,
                                        grads=None)
        # the fused optimizer reads gradients from .grad of the real Parameters
        self.optimizer._grad_list = self._grad_list  # type: ignore[method-assign]
        self._saved: Dict[int, Tuple[Optional[torch.Tensor], torch.Tensor]] = {}

    def _grad_list(self) -> List[torch.Tensor]:
        out = []
        for p in self.params:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            out.append(p.grad)
        return out

    def in_shape(self, mb_size: int):
        return (mb_size, *self._in)

    def out_shape(self, mb_size: int):
        return (mb_size, *self._out)

    def set_training(self, training: bool) -> None:
        self.module.train(training)

    def forward(self, mb, x, targets):
        if not self.is_first:
            x = x.detach().requires_grad_(True)
        y = self.module(x)
        if self.is_last:
            # scale * sum CE ; accumulate the un-scaled sum for logging
            loss = loss_ops.cross_entropy(y, targets, self._scale)
            self.loss_sum += loss.detach() / self._scale
            self._saved[mb] = (None if self.is_first else x, loss)
            return None
        self._saved[mb] = (None if self.is_first else x, y)
        return y.detach()

    def backward_input(self, mb, dy):
        x, y = self._saved.pop(mb)
        if self.is_last:
            y.backward()
        else:
            torch.autograd.backward(y, dy)
        return None if self.is_first else x.grad

    def backward_weight(self, mb) -> None:
        return None

    def enable_fused_dp(self, groups, topo, rank) -> None:
        """Replace NCCL all-reduce + optimizer launch by the one-kernel in-NVSwitch-domain
        all-reduce + SGD (``csrc/dp_allreduce_sgd.cu``).  SGD only; every replica calls this."""
        from ..ops.optim import FusedSGD
        from .symm import FusedDPSGD
        if not isinstance(self.optimizer, FusedSGD) or self.device.type != "cuda":
            return
        self._fused_dp = FusedDPSGD([p.data for p in self.params], self._grad_list,
                                    self.optimizer.momentum_buffers, self.optimizer.lr,
                                    self.optimizer.momentum, self.optimizer.weight_decay,
                                    topo.dp_group_ranks(rank), rank, groups.ctrl, self.device,
                                    timeout_s=getattr(self, "timeout_s", 300.0))

    def end_step(self) -> None:
        grad_scale = 1.0
        if getattr(self, "_fused_dp", None) is not None:
            self._fused_dp.step(grad_scale=1.0 / self.dp_size)
            self.optimizer.step_count += 1
            return
        if self.dp_group is not None and self.dp_size > 1:
            import torch.distributed as dist
            from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors
            grads = self._grad_list()
            flat = _flatten_dense_tensors(grads)
            dist.all_reduce(flat, group=self.dp_group)
            for g, f in zip(grads, _unflatten_dense_tensors(flat, grads)):
                g.copy_(f)
            grad_scale = 1.0 / self.dp_size
        self.optimizer.step(grad_scale=grad_scale, zero_grad=True)

    @torch.no_grad()
    def eval_forward(self, x, targets):
        y = self.module(x)
        if self.is_last:
            loss_ops.softmax_ce(y, targets, want_grad=False, want_correct=True,
                                loss_sum=self.loss_sum, correct=self.correct)
            return None
        return y

    def state_dict(self):
        return {k: v.detach().cpu().clone() for k, v in self.module.state_dict().items()}

    def load_state_dict(self, sd):
        self.module.load_state_dict(sd, strict=True)

    def checkpoint_part_indices(self, pp_idx):
        return sorted(self.part_prefixes) if self.part_prefixes else [pp_idx]

    def checkpoint_parts(self, pp_idx):
        if not self.part_prefixes:
            return {pp_idx: self.state_dict()}
        full = self.state_dict()
        return {idx: {k[len(pre):]: v for k, v in full.items() if k.startswith(pre)}
                for idx, pre in self.part_prefixes.items()}

    def load_checkpoint_parts(self, pp_idx, parts):
        if not self.part_prefixes:
            return self.load_state_dict(parts[pp_idx])
        merged = {}
        for idx, pre in self.part_prefixes.items():
            merged.update({pre + k: v for k, v in parts[idx].items()})
        self.load_state_dict(merged)
