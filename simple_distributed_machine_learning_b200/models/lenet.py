"""The reference model, split at the reference's split point.

Stage 0 = ``Network1`` without the RPC plumbing (/root/reference/simple_distributed.py:26-46):
conv1(1->10,k5) -> maxpool2 -> ReLU -> conv2(10->20,k5) -> Dropout2d -> maxpool2 -> ReLU ->
flatten [B,320].  Stage 1 = ``Network2`` (ref:60-80): fc1(320->50) -> ReLU -> dropout ->
fc2(50->10); the ``log_softmax`` (ref:79) is fused with the loss on this stage instead of
shipping log-probabilities back to rank 0 (ref:80, ref:111).

``state_dict()`` keys are exactly those of the reference classes (``conv1.weight``,
``conv1.bias``, ``conv2.weight``, ``conv2.bias`` / ``fc1.weight``, ``fc1.bias``,
``fc2.weight``, ``fc2.bias``), so checkpoints load into ``Network1`` / ``Network2``.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.linear import linear_f32


class LeNetStage0(nn.Module):
    """Convolutional half (rank 0 / "master" in the reference)."""

    out_features = 320

    def __init__(self, dropout: bool = True):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d() if dropout else nn.Identity()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        z1 = F.relu(F.max_pool2d(self.conv1(x), 2))
        z2 = self.conv2(z1)
        z3 = F.relu(F.max_pool2d(self.conv2_drop(z2), 2))
        return z3.reshape(-1, 320)


class LeNetStage1(nn.Module):
    """Fully-connected half (rank 1 / "worker1" in the reference); returns *logits*."""

    in_features = 320
    num_classes = 10

    def __init__(self, dropout: bool = True, ref_eval_dropout: bool = False, seed: int = 0):
        super().__init__()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)
        self.drop_p = 0.5 if dropout else 0.0
        # ref:75 calls F.dropout with its default training=True and Network2 is never put in
        # eval mode (ref:120 only reaches Network1's sub-modules), so the reference keeps
        # dropping activations at test time.  Off by default; opt-in for bit-for-bit behaviour.
        self.ref_eval_dropout = ref_eval_dropout
        self.seed = seed
        # device-resident RNG counter: advanced ON THE DEVICE once per forward call, so a
        # training step replayed from a CUDA graph still draws a fresh mask every time
        self.register_buffer("rng_counter", torch.zeros(1, dtype=torch.int64), persistent=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        p = self.drop_p if (self.training or self.ref_eval_dropout) else 0.0
        off = None
        if p > 0.0 and x.is_cuda:
            off = self.rng_counter.clone()          # value this call (and its backward) uses
            self.rng_counter.add_(1 << 24)          # disjoint counter range for the next call
        z4 = linear_f32(x, self.fc1.weight, self.fc1.bias, relu=True, drop_p=p, seed=self.seed,
                        offset_dev=off)
        return linear_f32(z4, self.fc2.weight, self.fc2.bias, relu=False)


class LeNetFull(nn.Module):
    """Single-process oracle: both halves in one module (used by parity tests and pp=1)."""

    def __init__(self, dropout: bool = True, ref_eval_dropout: bool = False, seed: int = 0):
        super().__init__()
        self.stage0 = LeNetStage0(dropout)
        self.stage1 = LeNetStage1(dropout, ref_eval_dropout, seed)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.stage1(self.stage0(x))


REPLICA_SEED_STRIDE = 1000003


def build_lenet_stage(stage_idx: int, pp: int, dropout: bool = True, ref_eval_dropout: bool = False,
                      seed: int = 0, replica: int = 0) -> nn.Module:
    """Module for pipeline stage ``stage_idx`` of ``pp`` (1 or 2).  Parameter init is seeded
    per stage so any (pp, dp) layout starts from identical weights; the DROPOUT stream is
    additionally keyed by the data-parallel ``replica`` (row i of two replicas must not share a mask)."""
    module = _build_lenet_stage(stage_idx, pp, dropout, ref_eval_dropout, seed)
    for m in module.modules():
        if isinstance(m, LeNetStage1):
            m.seed = seed + REPLICA_SEED_STRIDE * replica
    return module


def _build_lenet_stage(stage_idx: int, pp: int, dropout: bool, ref_eval_dropout: bool, seed: int) -> nn.Module:
    with torch.random.fork_rng(devices=[]):
        if pp == 1:
            torch.manual_seed(1000 + seed)
            s0 = LeNetStage0(dropout)
            torch.manual_seed(2000 + seed)
            s1 = LeNetStage1(dropout, ref_eval_dropout, seed)
            full = LeNetFull.__new__(LeNetFull)
            nn.Module.__init__(full)
            full.stage0, full.stage1 = s0, s1
            return full
        if stage_idx == 0:
            torch.manual_seed(1000 + seed)
            return LeNetStage0(dropout)
        if stage_idx == 1:
            torch.manual_seed(2000 + seed)
            return LeNetStage1(dropout, ref_eval_dropout, seed)
    raise ValueError("the reference model has exactly two stages")
