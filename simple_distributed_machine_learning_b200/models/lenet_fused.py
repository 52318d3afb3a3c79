"""The reference model on the fused kernels of ``csrc/lenet_fused.cu``.

Same parameters, same ``state_dict`` keys and the same arithmetic as
:mod:`.lenet` (``Network1`` / ``Network2`` of /root/reference/simple_distributed.py:26-83),
but a training step of a stage is one or two kernels instead of ~25 ATen launches:

  stage 0 : ``sdml_lenet_s0_fwd`` (conv-pool-relu x2 + Dropout2d, all in shared memory) and
            ``sdml_lenet_s0_bwd`` (their whole backward, per-sample, gradients by atomics);
  stage 1 : ``sdml_lenet_s1`` -- fc1+ReLU+dropout+fc2+log-softmax+NLL forward AND backward in
            one kernel per call (the gradient for stage 0 exists as soon as the forward of the
            last stage has run), plus the batched fc1 weight-gradient kernel.

Used on GPUs; the autograd modules in :mod:`.lenet` remain the CPU path and the oracle.
"""
from __future__ import annotations

from typing import Dict, List

import torch

from ..ops import _lib
from ..ops.optim import make_optimizer
from ..parallel.stage import Stage
from .lenet import build_lenet_stage


def _s():
    return torch.cuda.current_stream().cuda_stream


class LeNetFusedStage(Stage):
    """role = "full" (both halves on this rank), "conv" (stage 0) or "fc" (stage 1)."""

    act_dtype = torch.float32

    def __init__(self, pp_idx: int, pp: int, device: torch.device, dropout: bool = True,
                 ref_eval_dropout: bool = False, seed: int = 0, optimizer: str = "sgd", lr: float = 0.1,
                 momentum: float = 0.5, weight_decay: float = 0.0, replica: int = 0):
        super().__init__(pp_idx == 0, pp_idx == pp - 1, device)
        self.role = "full" if pp == 1 else ("conv" if pp_idx == 0 else "fc")
        self.module = build_lenet_stage(pp_idx, pp, dropout, ref_eval_dropout, seed).to(device)
        self.part_prefixes = {0: "stage0.", 1: "stage1."} if pp == 1 else None
        self.drop_p = 0.5 if dropout else 0.0
        self.ref_eval_dropout = ref_eval_dropout
        # dropout stream: same weights on every DP replica (seed), different masks (replica)
        from .lenet import REPLICA_SEED_STRIDE
        self.seed = seed + REPLICA_SEED_STRIDE * replica
        self.training = True
        conv = self.module.stage0 if pp == 1 else (self.module if self.role == "conv" else None)
        fc = self.module.stage1 if pp == 1 else (self.module if self.role == "fc" else None)
        self.conv, self.fc = conv, fc
        self.params: List[torch.Tensor] = [p.data for p in self.module.parameters()]
        self.grads: List[torch.Tensor] = [torch.zeros_like(p) for p in self.params]
        self._g = {p.data_ptr(): g for p, g in zip(self.params, self.grads)}
        self.optimizer = make_optimizer(optimizer, self.params, lr=lr, momentum=momentum,
                                        weight_decay=weight_decay, grads=self.grads)
        # Device-resident RNG positions, advanced once per step ON the device (CUDA-graph safe).
        # Each kernel advances the counter the OTHER kernel reads, so no extra launch is needed
        # when both halves live on this rank; a lone half uses one tiny add per step.
        self.rng_s0 = torch.zeros(1, dtype=torch.int64, device=device)   # read by the stage-0 forward
        self.rng_s1 = torch.zeros(1, dtype=torch.int64, device=device)   # read by the stage-1 kernel
        self.rng_step = self.rng_s0          # (name kept for the CUDA-graph state snapshot)
        self._wg_stream = torch.cuda.Stream(device=device) if self.role != "conv" else None
        self._wg_used = False
        self._saved: Dict[int, dict] = {}
        self._hold: List[dict] = []
        self._sizes: List[int] = []
        self._in_slots: Dict[int, torch.Tensor] = {}
        self._grad_slots: Dict[int, torch.Tensor] = {}

    # ---- helpers ----
    def _grad(self, p: torch.Tensor) -> torch.Tensor:
        return self._g[p.data_ptr()]

    def _grad_list(self) -> List[torch.Tensor]:
        return self.grads

    def in_shape(self, mb_size: int):
        return (mb_size, 1, 28, 28) if self.role != "fc" else (mb_size, 320)

    def out_shape(self, mb_size: int):
        return (mb_size, 320) if self.role == "conv" else (mb_size, 10)

    def set_training(self, training: bool) -> None:
        self.training = training
        self.module.train(training)

    # ---- static per-micro-batch receive slots (needed by the native executor, whose per-op CUDA
    # graphs bake buffer addresses in; off by default: the Python executor allocates per receive) ----
    static_slots = False
    # symmetric stage boundary (parallel/boundary.py): receive slots live in this rank's peer-mapped
    # region and the neighbour writes them directly
    supports_symm_boundary = True
    boundary = None
    # a pipeline half is a few ~10 us kernels on <= 8 CTAs per micro-batch: the executor may run its
    # forward and backward ops on two streams (parallel/pipeline.py)
    concurrent_fb = True

    def _slot(self, table: dict, mb: int, shape) -> torch.Tensor:
        t = table.get(mb)
        if t is None or tuple(t.shape) != tuple(shape):
            t = table[mb] = torch.empty(shape, dtype=torch.float32, device=self.device)
        return t

    def input_buffer(self, mb: int):
        if self.boundary is not None and self.role == "fc":
            return self.boundary.in_slot("act", mb, (320,))
        if not self.static_slots or self.role != "fc" or mb >= len(self._sizes):
            return None
        return self._slot(self._in_slots, mb, self.in_shape(self._sizes[mb]))

    def grad_buffer(self, mb: int):
        if self.boundary is not None and self.role == "conv":
            return self.boundary.in_slot("grad", mb, (320,))
        if not self.static_slots or self.role != "conv" or mb >= len(self._sizes):
            return None
        return self._slot(self._grad_slots, mb, self.out_shape(self._sizes[mb]))

    def begin_step(self, global_batch: int, sizes=None) -> None:
        self._scale = 1.0 / float(global_batch)
        self._single_mb = sizes is None or len(sizes) == 1
        self._sizes = list(sizes) if sizes is not None else [global_batch]
        if self.role != "full":
            self.loss_sum.zero_()            # "full": the stage-0 forward of micro-batch 0 clears it

    def mean_loss(self) -> torch.Tensor:
        return self.loss_sum                 # the training kernel accumulates loss * (1 / batch)

    # ---- kernels ----
    def _conv_fwd(self, x: torch.Tensor, mb: int, drop_p: float, housekeeping: bool = False) -> dict:
        B = x.shape[0]
        dev = self.device
        s = dict(x=x.contiguous(), z3=torch.empty(B, 320, device=dev), z1=torch.empty(B, 1440, device=dev),
                 idx1=torch.empty(B, 1440, dtype=torch.uint8, device=dev),
                 idx2=torch.empty(B, 320, dtype=torch.uint8, device=dev), mask=torch.empty(B, 20, device=dev))
        c = self.conv
        rc = _lib.cuda_lib().sdml_lenet_s0_fwd(
            s["x"].data_ptr(), c.conv1.weight.data_ptr(), c.conv1.bias.data_ptr(), c.conv2.weight.data_ptr(),
            c.conv2.bias.data_ptr(), B, s["z3"].data_ptr(), s["z1"].data_ptr(), s["idx1"].data_ptr(),
            s["idx2"].data_ptr(), s["mask"].data_ptr(), float(drop_p), int(self.seed),
            self.rng_s0.data_ptr(), int(mb) << 24,
            self.loss_sum.data_ptr() if housekeeping else None,
            self.rng_s1.data_ptr() if housekeeping else None, 1 << 32, _s())
        _lib.check(rc, "sdml_lenet_s0_fwd")
        return s

    def _conv_bwd(self, s: dict, dz3: torch.Tensor) -> None:
        c = self.conv
        rc = _lib.cuda_lib().sdml_lenet_s0_bwd(
            s["x"].data_ptr(), c.conv2.weight.data_ptr(), s["z3"].data_ptr(), dz3.contiguous().data_ptr(),
            s["z1"].data_ptr(), s["idx1"].data_ptr(), s["idx2"].data_ptr(), s["mask"].data_ptr(),
            s["x"].shape[0], self._grad(c.conv1.weight).data_ptr(), self._grad(c.conv1.bias).data_ptr(),
            self._grad(c.conv2.weight).data_ptr(), self._grad(c.conv2.bias).data_ptr(), _s())
        _lib.check(rc, "sdml_lenet_s0_bwd")

    def _fc(self, z3: torch.Tensor, targets: torch.Tensor, mb: int, train: bool, drop_p: float,
            housekeeping: bool = False):
        B = z3.shape[0]
        f = self.fc
        z3 = z3.contiguous()
        dz3 = torch.empty(B, 320, device=self.device) if train else None
        dz4 = torch.empty(B, 50, device=self.device) if train else None
        side = self._wg_stream if train else None
        if side is not None:
            z3.record_stream(side)
            dz4.record_stream(side)
            self._wg_used = True
        rc = _lib.cuda_lib().sdml_lenet_s1(
            z3.data_ptr(), targets.data_ptr(), f.fc1.weight.data_ptr(), f.fc1.bias.data_ptr(),
            f.fc2.weight.data_ptr(), f.fc2.bias.data_ptr(), B, int(train), float(self._scale), float(drop_p),
            int(self.seed) + 77, self.rng_s1.data_ptr(), int(mb) << 24, self.loss_sum.data_ptr(),
            float(self._scale) if train else 1.0,
            None if train else self.correct.data_ptr(),
            dz3.data_ptr() if train else None, dz4.data_ptr() if train else None,
            self._grad(f.fc1.weight).data_ptr() if train else None,
            self._grad(f.fc1.bias).data_ptr() if train else None,
            self._grad(f.fc2.weight).data_ptr() if train else None,
            self._grad(f.fc2.bias).data_ptr() if train else None,
            self.rng_s0.data_ptr() if housekeeping else None, 1 << 32, _s(),
            side.cuda_stream if side is not None else None)
        _lib.check(rc, "sdml_lenet_s1")
        return dz3

    # ---- schedule ops ----
    def forward(self, mb, x, targets):
        if self.role == "fc":
            self._saved[mb] = dict(dz3=self._fc(x, targets, mb, True, self.drop_p))
            return None
        hk = self.role == "full" and mb == 0
        s = self._conv_fwd(x, mb, self.drop_p, housekeeping=hk)
        if self.role == "full":
            s["dz3"] = self._fc(s["z3"], targets, mb, True, self.drop_p, housekeeping=hk)
            self._saved[mb] = s
            return None
        self._saved[mb] = s
        return s["z3"]

    def backward_input(self, mb, dy):
        s = self._saved.pop(mb)
        self._hold.append(s)      # allocated on the forward stream, read on the backward stream: free at end_step
        if self.role == "fc":
            return s["dz3"]
        self._conv_bwd(s, s["dz3"] if self.role == "full" else dy)
        return None

    def backward_weight(self, mb) -> None:
        return None

    def enable_fused_dp(self, groups, topo, rank) -> None:
        from ..ops.optim import FusedSGD
        from ..parallel.symm import FusedDPSGD
        if not isinstance(self.optimizer, FusedSGD):
            return
        opt, ranks = self.optimizer, topo.dp_group_ranks(rank)
        self._fused_dp = FusedDPSGD(self.params, self._grad_list, opt.momentum_buffers, opt.lr, opt.momentum,
                                    opt.weight_decay, ranks, rank, groups.ctrl, self.device,
                                    timeout_s=getattr(self, "timeout_s", 300.0))

    def join_side_streams(self) -> None:
        if self._wg_used:
            torch.cuda.current_stream(self.device).wait_stream(self._wg_stream)   # join the fc1 wgrad
            self._wg_used = False

    def end_step(self) -> None:
        self.join_side_streams()
        self._hold.clear()
        if getattr(self, "_fused_dp", None) is not None:
            self._fused_dp.step(grad_scale=1.0 / self.dp_size)
            self.optimizer.step_count += 1
        else:
            grad_scale = 1.0
            if self.dp_group is not None and self.dp_size > 1:
                import torch.distributed as dist
                from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors
                flat = _flatten_dense_tensors(self.grads)
                dist.all_reduce(flat, group=self.dp_group)
                for g, f in zip(self.grads, _unflatten_dense_tensors(flat, self.grads)):
                    g.copy_(f)
                grad_scale = 1.0 / self.dp_size
            self.optimizer.step(grad_scale=grad_scale, zero_grad=True)
        if self.role == "conv":
            self.rng_s0.add_(1 << 32)
        elif self.role == "fc":
            self.rng_s1.add_(1 << 32)

    @torch.no_grad()
    def eval_forward(self, x, targets):
        p1 = self.drop_p if self.ref_eval_dropout else 0.0   # ref:75 keeps stage-1 dropout on at eval
        if self.role == "fc":
            self._fc(x, targets, 0, False, p1)
            return None
        s = self._conv_fwd(x, 0, 0.0)
        if self.role == "full":
            self._fc(s["z3"], targets, 0, False, p1)
            return None
        return s["z3"]

    # ---- checkpoint (same files / keys as the autograd stages) ----
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
