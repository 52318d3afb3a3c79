"""Tensor-parallel wide-MLP stage (BASELINE.json config 5: pipeline x 2-way TP stages).

Layers come in pairs, Megatron style, with the activation sharded along the *batch* (row)
dimension between pairs, so the stage boundary carries ``rows/tp`` per TP rank:

  layer A (column-parallel)  y1_t = relu( all-gather_rows(x) . Wa_t^T + ba_t )     [rows, H/tp]
  layer B (row-parallel)     y2   = relu( reduce-scatter_rows( y1_t . Wb_t^T ) + bb ) [rows/tp, H]

Both collectives are fused into the GEMM kernels (``tp_ops.AllGatherGemm`` /
``GemmReduceScatter``: tcgen05 CTA-pair GEMM pulling peer rows over NVLink while it
computes, and TMA reduce-add of partial tiles into the row owner's memory).  The backward
pass uses the same two fused ops (all-gather(dz2) -> dgrad of B, dgrad of A ->
reduce-scatter), and the weight gradients reuse the already-gathered operands.

``backend="nccl"`` runs the identical math with ``all_gather_into_tensor`` /
``reduce_scatter_tensor`` + the plain GEMM -- the baseline the fused path is measured
against, and the path used on CPUs (gloo) by the tests.

``dtype="fp8"`` (BASELINE.json config 5 asks for block-scaled fp8 in the TP stages): every GEMM
of the stage -- forward, dgrad and wgrad of both layers of a pair -- runs on the MX-fp8
block-scaled kernels (``ops/fp8.py``); each gathered activation / gradient is quantised once
into a row-wise copy (forward / dgrad operand) and a transposed copy scaled along the rows
(wgrad operand).  In this mode the GEMM -> reduce-scatter half is in-kernel (the fp8 CTA-pair GEMM's
epilogue stores the partial rows into the owners' landing slots); the all-gather half is NCCL on bf16
rows by default.  An opt-in PUSH form exists (``tp_ops.Fp8AllGather``, ``SDML_TP_FP8_FUSED_AG=1``): every
rank quantises only its row shard and copies the fp8 bytes + scale atoms into its peers' operand
buffers over NVLink, flag-gated on the consumer's stream -- bit-identical, but measured slower than
NCCL + quantise (no overlap with the GEMM yet).  The tile-level AG -> GEMM overlap of the bf16 kernel
(gather warps inside the GEMM) exists for bf16 operands only.
"""
from __future__ import annotations

import math
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..models.mlp import _colsum, _gemm
from ..ops import loss as loss_ops
from ..ops.optim import make_optimizer
from .stage import Stage


class TPMLPStage(Stage):
    act_dtype = torch.bfloat16

    def __init__(self, stage_idx: int, pp: int, layers_total: int, hidden: int, classes: int,
                 device: torch.device, optimizer: str = "sgd", lr: float = 0.1, momentum: float = 0.5,
                 weight_decay: float = 0.0, seed: int = 0, tp: int = 2, tp_idx: int = 0,
                 dtype: str = "bf16", backend: str = "fused"):
        super().__init__(stage_idx == 0, stage_idx == pp - 1, device)
        assert layers_total % pp == 0 and (layers_total // pp) % 2 == 0, "TP stages hold layer pairs"
        self.stage_idx, self.pp, self.tp, self.tp_idx = stage_idx, pp, tp, tp_idx
        self.n_layers = layers_total // pp
        self.n_pairs = self.n_layers // 2
        self.hidden = hidden
        self.classes = classes or hidden
        assert hidden % tp == 0
        self.hs = hidden // tp
        self.backend = backend if device.type == "cuda" else "nccl"
        self.fp8 = dtype == "fp8" and os.environ.get("SDML_TP_FP8", "1") == "1"
        # fp8 stages: the GEMM -> reduce-scatter half runs in-kernel (block-scaled CTA-pair GEMM whose
        # epilogue TMA-stores the partial rows into the owners' landing slots, tp_ops.Fp8GemmReduceScatter);
        # the all-gather half is still NCCL (bf16 rows gathered, then quantised once per rank)
        self.fp8_fused_rs = False
        self.fp8_fused_ag = False
        if self.fp8:
            assert hidden % 128 == 0 and self.hs % 128 == 0 and self.classes % 128 == 0, \
                "fp8 TP stages need hidden, hidden/tp and classes to be multiples of 128"
            self.fp8_fused_rs = (self.backend == "fused" and tp > 1
                                 and os.environ.get("SDML_TP_FP8_FUSED_RS", "1") == "1")
            # all-gather half, OPT-IN (SDML_TP_FP8_FUSED_AG=1): every rank quantises only its shard and
            # pushes the fp8 rows + scale atoms into its peers' operand buffers (tp_ops.Fp8AllGather)
            # instead of an NCCL gather of bf16 rows followed by a quantiser pass over the whole operand.
            # Bit-identical (tests/test_tp_gpu.py) but MEASURED SLOWER at tp2 / hidden 16384 (61.9 vs
            # 58.2 ms per step): the push is not overlapped with the GEMM, and the column range of the
            # transposed copy travels as 16 384 segments of 1 KiB.  Default: NCCL gather + fused reduce-scatter.
            self.fp8_fused_ag = (self.backend == "fused" and tp > 1
                                 and os.environ.get("SDML_TP_FP8_FUSED_AG", "0") == "1")
            self.backend = "nccl"       # the fused AG -> GEMM kernel is bf16-only
        self.tp_group = None
        first_layer = stage_idx * self.n_layers
        # full-model init (same generators as MLPStage), then slice this rank's shard
        self.wa_m, self.ba_m, self.wb_m, self.bb_m = [], [], [], []
        self.wa, self.wb = [], []
        self.dwa, self.dba, self.dwb, self.dbb = [], [], [], []
        self.out_dims: List[int] = []
        for j in range(self.n_pairs):
            la, lb = first_layer + 2 * j, first_layer + 2 * j + 1
            out_b = self.classes if lb == layers_total - 1 else hidden
            self.out_dims.append(out_b)
            wa_full, ba_full = self._init_layer(la, hidden, hidden, seed)
            wb_full, bb_full = self._init_layer(lb, out_b, hidden, seed)
            sl = slice(tp_idx * self.hs, (tp_idx + 1) * self.hs)
            wa_t = wa_full[sl, :].contiguous().to(device)          # [H/tp, H]
            ba_t = ba_full[sl].contiguous().to(device)
            wb_t = wb_full[:, sl].contiguous().to(device)          # [out, H/tp]
            self.wa_m.append(wa_t); self.ba_m.append(ba_t)
            self.wb_m.append(wb_t); self.bb_m.append(bb_full.to(device))
            self.wa.append(wa_t.to(torch.bfloat16)); self.wb.append(wb_t.to(torch.bfloat16))
            self.dwa.append(torch.zeros_like(wa_t)); self.dba.append(torch.zeros_like(ba_t))
            self.dwb.append(torch.zeros_like(wb_t)); self.dbb.append(torch.zeros(out_b, device=device))
        self.last_is_logits = self.is_last
        params = self.wa_m + self.wb_m + self.ba_m + self.bb_m
        grads = self.dwa + self.dwb + self.dba + self.dbb
        lowp = self.wa + self.wb + [None] * (2 * self.n_pairs)
        self.optimizer = make_optimizer(optimizer, params, lr=lr, momentum=momentum,
                                        weight_decay=weight_decay, grads=grads, lowp=lowp)
        self.w_group = 0
        self._rows = 0          # rows of one micro-batch (full, before TP sharding)
        self._n_mb = 0
        self._pending_w: List[int] = []
        self._w_done_any = False
        self._ctx = None        # (topo ranks, my rank, ctrl group) for symmetric allocations

    @staticmethod
    def _init_layer(global_layer: int, fo: int, fi: int, seed: int):
        g = torch.Generator().manual_seed(31337 + 101 * global_layer + seed)
        bound = math.sqrt(6.0 / fi)
        w = (torch.rand(fo, fi, generator=g) * 2 - 1) * bound
        b = (torch.rand(fo, generator=g) * 2 - 1) / math.sqrt(fi)
        return w, b

    def attach_tp(self, tp_group, tp_ranks: List[int], my_rank: int, ctrl_group) -> None:
        self.tp_group = tp_group
        self._ctx = (list(tp_ranks), my_rank, ctrl_group)

    # ---- shapes: what crosses the stage boundary is this rank's row shard ----
    def in_shape(self, mb_size: int):
        # the executor counts in this rank's rows (already rows/tp per micro-batch)
        return (mb_size, self.hidden)

    def out_shape(self, mb_size: int):
        return (mb_size, self.out_dims[-1])

    # ---- buffers ----
    def _alloc(self, rows: int, n_mb: int) -> None:
        from . import tp_ops
        dev, H, hs = self.device, self.hidden, self.hs
        self._rows, self._n_mb = rows, n_mb
        rs = rows // self.tp
        B = rows * n_mb
        self.Y1 = [torch.empty(B, hs, dtype=torch.bfloat16, device=dev) for _ in range(self.n_pairs)]
        self.DZ1 = [torch.empty(B, hs, dtype=torch.bfloat16, device=dev) for _ in range(self.n_pairs)]
        self.OUT = torch.empty(rs * n_mb, self.out_dims[-1], dtype=torch.bfloat16, device=dev)
        if self.backend == "fused":
            ranks, me, ctrl = self._ctx
            self.agX = [tp_ops.AllGatherGemm(ranks, me, ctrl, rows, H, dev, slots=n_mb) for _ in range(self.n_pairs)]
            self.agD = [tp_ops.AllGatherGemm(ranks, me, ctrl, rows, self.out_dims[j], dev, slots=n_mb)
                        for j in range(self.n_pairs)]
            self.rsF = [tp_ops.GemmReduceScatter(ranks, me, ctrl, rows, self.out_dims[j], dev)
                        for j in range(self.n_pairs)]
            self.rsB = [tp_ops.GemmReduceScatter(ranks, me, ctrl, rows, H, dev) for _ in range(self.n_pairs)]
            self.Xfull = [a.all for a in self.agX]
            self.DZ2full = [a.all for a in self.agD]
        else:
            self.Xfull = [torch.empty(B, H, dtype=torch.bfloat16, device=dev) for _ in range(self.n_pairs)]
            self.DZ2full = [torch.empty(B, self.out_dims[j], dtype=torch.bfloat16, device=dev)
                            for j in range(self.n_pairs)]
        if self.fp8_fused_rs:
            ranks, me, ctrl = self._ctx
            self.rsF = [tp_ops.Fp8GemmReduceScatter(ranks, me, ctrl, rows, self.out_dims[j], dev)
                        for j in range(self.n_pairs)]
            self.rsB = [tp_ops.Fp8GemmReduceScatter(ranks, me, ctrl, rows, H, dev) for _ in range(self.n_pairs)]
        if self.fp8:
            from ..ops.fp8 import MXFP8Transposed
            assert rows % 128 == 0, "fp8 TP stages need micro-batches that are multiples of 128 rows"
            # transposed (scaled along the rows) copies of a whole step, K-sliced by the wgrad GEMMs
            if self.fp8_fused_ag:
                ranks, me, ctrl = self._ctx
                tmo = getattr(self, "timeout_s", 300.0)
                self.agXq = [tp_ops.Fp8AllGather(ranks, me, ctrl, rows, H, dev, slots=n_mb, timeout_s=tmo)
                             for _ in range(self.n_pairs)]
                self.agDq = [tp_ops.Fp8AllGather(ranks, me, ctrl, rows, self.out_dims[j], dev, slots=n_mb, timeout_s=tmo)
                             for j in range(self.n_pairs)]
                # the step-wide transposed operands of the wgrad GEMMs live in the gather buffers
                self.XT = [a.transposed for a in self.agXq]
                self.DZ2T = [a.transposed for a in self.agDq]
            else:
                self.XT = [MXFP8Transposed(B, H, dev) for _ in range(self.n_pairs)]
                self.DZ2T = [MXFP8Transposed(B, self.out_dims[j], dev) for j in range(self.n_pairs)]
            self.Y1T = [MXFP8Transposed(B, hs, dev) for _ in range(self.n_pairs)]
            self.DZ1T = [MXFP8Transposed(B, hs, dev) for _ in range(self.n_pairs)]
            if getattr(self, "_waq", None) is None:
                self._requantize_weights()

    def begin_step(self, global_batch: int, sizes: Optional[List[int]] = None) -> None:
        # `global_batch` counts this rank's rows only (rows/tp per micro-batch): the loss is
        # averaged over the full batch, so rescale.
        sizes = list(sizes) if sizes is not None else [global_batch]
        assert len(set(sizes)) == 1, "TP stages need equal micro-batches"
        rows = sizes[0] * self.tp
        super().begin_step(global_batch * self.tp)
        if rows != self._rows or len(sizes) != self._n_mb:
            self._alloc(rows, len(sizes))
        self._pending_w = []
        self._w_done_any = False

    def _mb(self, buf: torch.Tensor, mb: int, rows: int) -> torch.Tensor:
        return buf[mb * rows:(mb + 1) * rows]

    def _shard(self, buf: torch.Tensor, mb: int) -> torch.Tensor:
        lo = mb * self._rows + self.tp_idx * (self._rows // self.tp)
        return buf[lo:lo + self._rows // self.tp]

    def input_buffer(self, mb: int):
        return self._shard(self.Xfull[0], mb)

    def grad_buffer(self, mb: int):
        return self._shard(self.DZ2full[-1], mb)

    # ---- fp8: weights quantised in place after every optimizer step ----
    def _requantize_weights(self) -> None:
        from ..ops.fp8 import quantize_mxfp8_dual
        if getattr(self, "_waq", None) is None:
            n = self.n_pairs
            self._waq, self._waqT, self._wbq, self._wbqT = [None] * n, [None] * n, [None] * n, [None] * n
        for j in range(self.n_pairs):
            # Wa_t [hs, H]: row-wise for the forward (K = H), transposed [H, hs] for dgrad (K = hs)
            self._waq[j], self._waqT[j] = quantize_mxfp8_dual(self.wa[j], want_col=True, row_out=self._waq[j],
                                                              col_out=self._waqT[j])
            # Wb_t [out, hs]: row-wise for the forward (K = hs), transposed [hs, out] for dgrad (K = out)
            self._wbq[j], self._wbqT[j] = quantize_mxfp8_dual(self.wb[j], want_col=True, row_out=self._wbq[j],
                                                              col_out=self._wbqT[j])

    # ---- the two fused ops (or their NCCL baseline) ----
    def _ag_gemm(self, which: str, j: int, mb: int, W, C, *, b_mn=0, bias=None, aux=None, relu=False):
        full_buf = self.Xfull[j] if which == "x" else self.DZ2full[j]
        if self.backend == "fused":
            op = self.agX[j] if which == "x" else self.agD[j]
            op.run(W, C, b_mn=b_mn, bias=bias, aux=aux, relu=relu, slot=mb)
            return
        if self.fp8_fused_ag:
            from ..ops.fp8 import gemm_mxfp8
            op = self.agXq[j] if which == "x" else self.agDq[j]
            q = op.gather(mb, self._shard(full_buf, mb))      # my shard quantised + pushed, peers' rows awaited
            gemm_mxfp8(q, self._waq[j] if which == "x" else self._wbqT[j], C, bias=bias, aux=aux, relu=relu)
            return
        full = self._mb(full_buf, mb, self._rows)
        shard = self._shard(full_buf, mb).clone()
        if full.is_cuda:
            dist.all_gather_into_tensor(full, shard, group=self.tp_group)
        else:
            parts = [torch.empty_like(shard) for _ in range(self.tp)]
            dist.all_gather(parts, shard, group=self.tp_group)
            full.copy_(torch.cat(parts, 0))
        if self.fp8:
            # one pass over the gathered operand: row-wise copy for this GEMM, transposed copy for
            # the weight-gradient GEMM that contracts over the rows
            from ..ops.fp8 import gemm_mxfp8, quantize_mxfp8_dual
            if which == "x":
                q, _ = quantize_mxfp8_dual(full, col_into=self.XT[j], row_offset=mb * self._rows)
                gemm_mxfp8(q, self._waq[j], C, bias=bias, aux=aux, relu=relu)       # x . Wa_t^T
            else:
                q, _ = quantize_mxfp8_dual(full, col_into=self.DZ2T[j], row_offset=mb * self._rows)
                gemm_mxfp8(q, self._wbqT[j], C, bias=bias, aux=aux, relu=relu)      # dz2 . Wb_t
            return
        _gemm(full, W, C, b_mn=b_mn, bias=bias, aux=aux, relu=relu)

    def _gemm_rs(self, which: str, j: int, A, W, out, *, b_mn=0, bias=None, aux=None, relu=False):
        if self.backend == "fused":
            op = self.rsF[j] if which == "f" else self.rsB[j]
            par = op.run(A, W, b_mn=b_mn)
            op.finalize(par, out, bias=bias, aux=aux, relu=relu)
            return
        N = W.shape[1] if b_mn else W.shape[0]
        if self.fp8_fused_rs:
            from ..ops.fp8 import quantize_mxfp8_dual
            op = self.rsF[j] if which == "f" else self.rsB[j]
            if which == "f":    # y1_t . Wb_t^T ; y1_t^T kept for dWb
                q, _ = quantize_mxfp8_dual(A, col_into=self.Y1T[j], row_offset=self._cur_mb * self._rows)
                par = op.run(q, self._wbq[j])
            else:               # dz1 . Wa_t ; dz1^T kept for dWa
                q, _ = quantize_mxfp8_dual(A, col_into=self.DZ1T[j], row_offset=self._cur_mb * self._rows)
                par = op.run(q, self._waqT[j])
            op.finalize(par, out, bias=bias, aux=aux, relu=relu)
            return
        partial = torch.empty(A.shape[0], N, dtype=torch.float32, device=A.device)
        if self.fp8:
            from ..ops.fp8 import gemm_mxfp8, quantize_mxfp8_dual
            if which == "f":    # y1_t . Wb_t^T ; y1_t^T kept for dWb
                q, _ = quantize_mxfp8_dual(A, col_into=self.Y1T[j], row_offset=self._cur_mb * self._rows)
                gemm_mxfp8(q, self._wbq[j], partial)
            else:               # dz1 . Wa_t ; dz1^T kept for dWa
                q, _ = quantize_mxfp8_dual(A, col_into=self.DZ1T[j], row_offset=self._cur_mb * self._rows)
                gemm_mxfp8(q, self._waqT[j], partial)
        else:
            _gemm(A, W, partial, b_mn=b_mn)
        rs = A.shape[0] // self.tp
        if partial.is_cuda:
            red = torch.empty(rs, N, dtype=torch.float32, device=A.device)
            dist.reduce_scatter_tensor(red, partial, group=self.tp_group)
        else:
            dist.all_reduce(partial, group=self.tp_group)
            red = partial[self.tp_idx * rs:(self.tp_idx + 1) * rs]
        if bias is not None:
            red = red + bias
        if relu:
            red = torch.relu(red)
        if aux is not None:
            red = red * (aux.float() > 0)
        out.copy_(red)

    # ---- schedule ops ----
    def forward(self, mb, x, targets):
        self._cur_mb = mb
        rs = self._rows // self.tp
        x0 = self._shard(self.Xfull[0], mb)
        if x.data_ptr() != x0.data_ptr():
            x0.copy_(x)
        for j in range(self.n_pairs):
            y1 = self._mb(self.Y1[j], mb, self._rows)
            self._ag_gemm("x", j, mb, self.wa[j], y1, bias=self.ba_m[j], relu=True)
            last = j == self.n_pairs - 1
            out = (self.OUT[mb * rs:(mb + 1) * rs] if last else self._shard(self.Xfull[j + 1], mb))
            self._gemm_rs("f", j, y1, self.wb[j], out, bias=self.bb_m[j],
                          relu=not (last and self.last_is_logits))
        if self.is_last:
            loss_ops.softmax_ce(out, targets, grad_scale=self._scale, want_grad=True,
                                loss_sum=self.loss_sum, dlogits=self._shard(self.DZ2full[-1], mb))
            return None
        return out

    def backward_input(self, mb, dy):
        self._cur_mb = mb
        dz2 = self._shard(self.DZ2full[-1], mb)
        if not self.is_last and dy.data_ptr() != dz2.data_ptr():
            dz2.copy_(dy)
        dx = None
        for j in range(self.n_pairs - 1, -1, -1):
            y1 = self._mb(self.Y1[j], mb, self._rows)
            dz1 = self._mb(self.DZ1[j], mb, self._rows)
            # dz1 = (all-gather(dz2) . Wb_t) * (y1 > 0)
            self._ag_gemm("d", j, mb, self.wb[j], dz1, b_mn=1, aux=y1)
            if j == 0 and self.is_first:
                if self.fp8:    # no dgrad below the first layer, but dWa still needs dz1^T
                    from ..ops.fp8 import quantize_mxfp8_dual
                    quantize_mxfp8_dual(dz1, want_row=False, col_into=self.DZ1T[j], row_offset=mb * self._rows)
                return None
            # dx_shard = reduce-scatter(dz1 . Wa_t) * (x_shard > 0)
            x_shard = self._shard(self.Xfull[j], mb)
            if j > 0:
                dst = self._shard(self.DZ2full[j - 1], mb)
            else:
                dst = torch.empty(self._rows // self.tp, self.hidden, dtype=torch.bfloat16, device=self.device)
            self._gemm_rs("b", j, dz1, self.wa[j], dst, b_mn=1, aux=x_shard)
            dx = dst
        return dx

    def _flush_w(self) -> None:
        if not self._pending_w:
            return
        lo, hi = self._pending_w[0] * self._rows, (self._pending_w[-1] + 1) * self._rows
        acc = self._w_done_any
        for j in range(self.n_pairs):
            dz2, y1 = self.DZ2full[j][lo:hi], self.Y1[j][lo:hi]
            dz1, x = self.DZ1[j][lo:hi], self.Xfull[j][lo:hi]
            if self.fp8:
                from ..ops.fp8 import gemm_mxfp8
                gemm_mxfp8(self.DZ2T[j].rows_slice(lo, hi), self.Y1T[j].rows_slice(lo, hi), self.dwb[j],
                           accumulate=acc)
                gemm_mxfp8(self.DZ1T[j].rows_slice(lo, hi), self.XT[j].rows_slice(lo, hi), self.dwa[j],
                           accumulate=acc)
            else:
                _gemm(dz2, y1, self.dwb[j], a_mn=1, b_mn=1, accumulate=acc)   # dWb_t = dz2^T . y1_t
                _gemm(dz1, x, self.dwa[j], a_mn=1, b_mn=1, accumulate=acc)    # dWa_t = dz1_t^T . x
            _colsum(dz2, self.dbb[j], accumulate=acc)
            _colsum(dz1, self.dba[j], accumulate=acc)
        self._w_done_any = True
        self._pending_w = []

    def flush_weight_grads(self) -> None:
        self._flush_w()

    def backward_weight(self, mb) -> None:
        if self._pending_w and mb != self._pending_w[-1] + 1:
            self._flush_w()
        self._pending_w.append(mb)
        if self.w_group and len(self._pending_w) >= self.w_group:
            self._flush_w()

    def end_step(self) -> None:
        self._flush_w()
        if self.fp8_fused_ag:
            for op in self.agXq + self.agDq:       # every reader of this step's operands is enqueued
                op.end_step()
        self.optimizer.step(grad_scale=1.0, zero_grad=False)
        if self.fp8:
            self._requantize_weights()

    def check_peers(self) -> None:
        """Raise if an in-kernel wait for a tensor-parallel peer has given up (pinned host words)."""
        if self.fp8_fused_ag and getattr(self, "agXq", None):
            for op in self.agXq + self.agDq:
                op.check()

    def mean_loss(self) -> torch.Tensor:
        # every TP rank holds the loss of its own rows; the caller may all-reduce for logging
        return self.loss_sum * self._scale

    # ---- evaluation ----
    @torch.no_grad()
    def eval_forward(self, x, targets):
        """Forward-only pass of one test batch (ref:119-132 runs the model under no_grad).

        Not row-sharded: every TP rank sees the whole batch (any size, ragged last batch
        included), computes its column shard of layer A and its partial sum of layer B, and one
        all-reduce per pair completes the row-parallel layer.  Evaluation is a few batches per
        epoch; the fused collectives are kept for the training path."""
        h = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
        n = h.shape[0]
        for j in range(self.n_pairs):
            y1 = torch.empty(n, self.hs, dtype=torch.bfloat16, device=self.device)
            _gemm(h, self.wa[j], y1, bias=self.ba_m[j], relu=True)
            part = torch.empty(n, self.out_dims[j], dtype=torch.float32, device=self.device)
            _gemm(y1, self.wb[j], part)
            if self.tp_group is not None:
                dist.all_reduce(part, group=self.tp_group)
            part += self.bb_m[j]
            if not (j == self.n_pairs - 1 and self.last_is_logits):
                part.relu_()
            h = part.to(torch.bfloat16)
        if self.is_last:
            loss_ops.softmax_ce(h, targets, want_grad=False, want_correct=True,
                                loss_sum=self.loss_sum, correct=self.correct)
            return None
        return h

    # ---- checkpoint: this rank's shards under nn.Module-style names ----
    def state_dict(self):
        sd = {}
        for j in range(self.n_pairs):
            sd[f"layers.{2 * j}.weight"] = self.wa_m[j].detach().cpu().clone()
            sd[f"layers.{2 * j}.bias"] = self.ba_m[j].detach().cpu().clone()
            sd[f"layers.{2 * j + 1}.weight"] = self.wb_m[j].detach().cpu().clone()
            sd[f"layers.{2 * j + 1}.bias"] = self.bb_m[j].detach().cpu().clone()
        return sd

    def load_state_dict(self, sd):
        for j in range(self.n_pairs):
            self.wa_m[j].copy_(sd[f"layers.{2 * j}.weight"]); self.ba_m[j].copy_(sd[f"layers.{2 * j}.bias"])
            self.wb_m[j].copy_(sd[f"layers.{2 * j + 1}.weight"]); self.bb_m[j].copy_(sd[f"layers.{2 * j + 1}.bias"])
            self.wa[j].copy_(self.wa_m[j]); self.wb[j].copy_(self.wb_m[j])
        if getattr(self, "_waq", None) is not None:
            self._requantize_weights()       # fp8 copies follow the loaded weights (in place)

    def flops_per_sample(self) -> int:
        """Model FLOPs per *full-batch* sample executed by this rank (fwd + dgrad + wgrad)."""
        total = 0
        for j in range(self.n_pairs):
            a = 2 * self.hs * self.hidden
            b = 2 * self.out_dims[j] * self.hs
            total += 3 * b + (2 if (j == 0 and self.is_first) else 3) * a
        return total
