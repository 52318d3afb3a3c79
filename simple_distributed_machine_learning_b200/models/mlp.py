"""Wide-MLP pipeline stages with explicit forward / dgrad / wgrad on the tcgen05 GEMM.

This is the model family of BASELINE.json configs 3-5 (N-stage MLP, hidden 8192 / 16384):
every layer is Linear + bias + ReLU -- the same op as the reference's fc1
(/root/reference/simple_distributed.py:63, :75) scaled up -- except the very last layer,
which produces the logits for the fused softmax cross-entropy.

Per layer and micro-batch (bf16 activations, fp32 master weights + bf16 compute copy):

  forward   y  = relu(x . W^T + b)              one GEMM, bias+ReLU in the epilogue
  dgrad     dz_prev = (dz . W) * (x > 0)        one GEMM, ReLU-backward mask in the epilogue
  wgrad     dW += dz^T . x ;  db += colsum(dz)  one GEMM accumulating in fp32 + a column sum

The gradient that crosses a stage boundary is therefore already ``dz`` of the previous
stage's last layer (the receiving side needs no elementwise pass), and dgrad / wgrad are
separately schedulable (zero-bubble schedules).
"""
from __future__ import annotations

import math
import os
from typing import List, Optional

import torch

from ..ops import gemm as gemm_ops
from ..ops import loss as loss_ops
from ..ops import _lib
from ..ops.optim import make_optimizer
from ..parallel.stage import Stage


# SDML_GEMM_BACKEND=cublas routes every GEMM of the MLP stages through torch.matmul (cuBLAS) with
# un-fused bias / ReLU / mask / accumulate ops: the "plain PyTorch + cuBLAS" baseline of
# SURVEY.md section 6.3 that the tcgen05 kernels are measured against.  Never the default.
_CUBLAS_BASELINE = os.environ.get("SDML_GEMM_BACKEND", "") == "cublas"


def _gemm_cublas(A, B, C, a_mn, b_mn, bias, aux, relu, accumulate, alpha):
    a = A.t() if a_mn else A
    b = B if b_mn else B.t()
    r = torch.matmul(a, b)
    if alpha != 1.0:
        r = r * alpha
    if bias is not None:
        r = r + bias.to(r.dtype)
    if relu:
        r = torch.relu(r)
    if aux is not None:
        r = r * (aux > 0)
    if accumulate:
        C += r
    else:
        C.copy_(r)
    return C


def _gemm(A, B, C, *, a_mn=0, b_mn=0, bias=None, aux=None, relu=False, accumulate=False, alpha=1.0):
    """tcgen05 kernel on CUDA; fp32 PyTorch oracle on CPU (plumbing tests)."""
    if A.is_cuda and _CUBLAS_BASELINE:
        return _gemm_cublas(A, B, C, a_mn, b_mn, bias, aux, relu, accumulate, alpha)
    if A.is_cuda:
        return gemm_ops.gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bias=bias, aux=aux, relu=relu,
                                  accumulate=accumulate, alpha=alpha)
    a = A.float().t() if a_mn else A.float()
    b = B.float().t() if b_mn else B.float()
    r = alpha * (a @ b.t())
    if bias is not None:
        r = r + bias
    if relu:
        r = torch.relu(r)
    if aux is not None:
        r = r * (aux.float() > 0)
    if accumulate:
        C += r
    else:
        C.copy_(r)
    return C


def _colsum(x: torch.Tensor, out: torch.Tensor, accumulate: bool = True) -> None:
    if x.is_cuda:
        rc = _lib.cuda_lib().sdml_colsum_bf16(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1],
                                              out.data_ptr(), int(accumulate),
                                              torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_colsum_bf16")
    else:
        s = x.float().sum(dim=0)
        if accumulate:
            out += s
        else:
            out.copy_(s)


class MLPStage(Stage):
    """``n_layers`` consecutive Linear(+ReLU) layers owned by one pipeline rank."""

    act_dtype = torch.bfloat16
    # Symmetric stage boundary (parallel/boundary.py): the receive buffers of a step -- X[0] and
    # dZ[-1] -- are carved out of this rank's peer-mapped region, and the GEMMs that produce what
    # crosses a boundary (last layer forward, first layer dgrad) write their output tiles straight
    # into the NEIGHBOUR's region through the TMA-store epilogue: the transfer is the epilogue.
    supports_symm_boundary = True
    boundary = None

    def __init__(self, stage_idx: int, pp: int, layers_total: int, hidden: int, classes: int,
                 device: torch.device, optimizer: str = "sgd", lr: float = 0.1,
                 momentum: float = 0.5, weight_decay: float = 0.0, seed: int = 0,
                 in_dim: Optional[int] = None, dtype: str = "bf16"):
        super().__init__(stage_idx == 0, stage_idx == pp - 1, device)
        # dtype "fp8": GEMMs run on the MX-fp8 block-scaled tensor-core path.  Forward: activations
        # quantised per call, weights re-quantised after every optimizer step.  Backward
        # (``fp8_backward``): MX scales must run along the contraction dimension, so every
        # activation / gradient / weight also gets a TRANSPOSED copy scaled along its rows (one
        # dual-output quantiser pass, ``sdml_quant_mxfp8_dual``) -- dgrad contracts dz with W^T,
        # wgrad contracts dz^T with x^T over the rows of a micro-batch group.  Layers or batch
        # shapes that are not multiples of 128 fall back to the bf16 kernels.
        self.fp8_forward = dtype == "fp8"      # (on CPUs: ops/fp8.py's emulation of the same arithmetic)
        self.fp8_backward = self.fp8_forward and os.environ.get("SDML_FP8_BWD", "1") == "1"
        assert layers_total % pp == 0
        self.stage_idx, self.pp = stage_idx, pp
        self.n_layers = layers_total // pp
        self.hidden = hidden
        self.classes = classes or hidden
        self.in_dim = in_dim or hidden
        first_layer = stage_idx * self.n_layers
        self.dims: List[tuple] = []
        for l in range(self.n_layers):
            gl = first_layer + l
            fan_in = self.in_dim if gl == 0 else hidden
            fan_out = self.classes if gl == layers_total - 1 else hidden
            self.dims.append((fan_out, fan_in))
        self.last_layer_is_logits = self.is_last
        # parameters: fp32 master + bf16 compute copy, deterministic per *global* layer index
        self.w_master: List[torch.Tensor] = []
        self.b_master: List[torch.Tensor] = []
        self.w: List[torch.Tensor] = []
        self.dw: List[torch.Tensor] = []
        self.db: List[torch.Tensor] = []
        for l, (fo, fi) in enumerate(self.dims):
            g = torch.Generator().manual_seed(31337 + 101 * (first_layer + l) + seed)
            bound = math.sqrt(6.0 / fi)          # He-uniform: keeps ReLU activations O(1)
            wm = (torch.rand(fo, fi, generator=g) * 2 - 1) * bound
            bm = (torch.rand(fo, generator=g) * 2 - 1) / math.sqrt(fi)
            self.w_master.append(wm.to(device))
            self.b_master.append(bm.to(device))
            self.w.append(wm.to(device=device, dtype=torch.bfloat16))
            self.dw.append(torch.zeros(fo, fi, dtype=torch.float32, device=device))
            self.db.append(torch.zeros(fo, dtype=torch.float32, device=device))
        params = self.w_master + self.b_master
        grads = self.dw + self.db
        lowp = self.w + [None] * len(self.b_master)
        self.optimizer = make_optimizer(optimizer, params, lr=lr, momentum=momentum,
                                        weight_decay=weight_decay, grads=grads, lowp=lowp)
        # Activations and pre-activation gradients of a whole step live in per-layer row-stacked
        # buffers: micro-batch i owns rows [off_i, off_{i+1}).  A layer writes its output straight
        # into the next layer's input slice, receives land directly in the slice, and -- the
        # point -- the weight gradient of a *group* of consecutive micro-batches is ONE GEMM over
        # their contiguous rows (K = rows of the group) instead of one fp32 read-modify-write of
        # the whole dW per micro-batch.
        # micro-batches per wgrad GEMM; 0 = all of the step at once (best without a pipeline: no
        # fp32 accumulate traffic); pipelines want small groups so W ops still fill bubbles
        self.w_group = int(os.environ.get("SDML_W_GROUP", 0 if pp == 1 else 2))
        self._rows = 0
        self._bufsets = {}                   # row count -> buffer set (see begin_step)
        self._fp8_fallback_logged = False
        self._offs: List[int] = [0]
        self._X: List[torch.Tensor] = []     # layer inputs  [B, in_l]
        self._Y: Optional[torch.Tensor] = None  # stage output [B, out_last]
        self._dZ: List[torch.Tensor] = []    # dz of every layer [B, out_l]
        self._pending_w: List[int] = []
        self._w_done_any = False
        self._fp8_bwd_ok = [False] * self.n_layers   # per layer: dgrad / wgrad on the fp8 kernels this step
        self._Xq_slots, self._dZq_slots = {}, {}
        self._producer_quant = False
        self._xq_ready, self._dzq_ready = set(), set()

    # ---- shapes ----
    def in_shape(self, mb_size: int):
        return (mb_size, self.dims[0][1])

    def out_shape(self, mb_size: int):
        return (mb_size, self.dims[-1][0])

    # ---- step set-up ----
    def begin_step(self, global_batch: int, sizes: Optional[List[int]] = None) -> None:
        super().begin_step(global_batch)
        sizes = list(sizes) if sizes is not None else [global_batch]
        offs = [0]
        for n in sizes:
            offs.append(offs[-1] + n)
        self._offs = offs
        if offs[-1] != self._rows:
            # One buffer set PER row count, kept for the life of the stage: a CUDA graph captured for a
            # batch size has these addresses baked in, so a ragged last batch of an epoch must not free
            # (and let the allocator recycle) the buffers the full-size graph replays into.
            if self._rows:
                self._bufsets[self._rows] = (self._X, self._Y, self._dZ, self._fp8_bwd_ok,
                                             getattr(self, "_XqT", None), getattr(self, "_dZqT", None),
                                             self._Xq_slots, self._dZq_slots)
            self._rows = offs[-1]
            cached = self._bufsets.get(self._rows)
            if cached is not None:
                (self._X, self._Y, self._dZ, self._fp8_bwd_ok, self._XqT, self._dZqT,
                 self._Xq_slots, self._dZq_slots) = cached
            else:
                self._alloc_buffers(sizes)
        elif self.fp8_backward and any(self._fp8_bwd_ok) and not all(n % 128 == 0 for n in sizes):
            self._fp8_bwd_ok = [False] * self.n_layers
        # producer-side quantisation: the GEMM that writes an activation / gradient also emits its fp8
        # copies (from its epilogue when ops.fp8._FUSED_QOUT, else by a pass right after it); the
        # consumer then skips its own quantiser pass.  (layer, micro-batch) pairs filled this step:
        from ..ops import fp8 as _fp8
        self._producer_quant = bool(_fp8._FUSED_QOUT) and any(self._fp8_bwd_ok)
        self._xq_ready, self._dzq_ready = set(), set()
        self._pending_w = []
        self._w_done_any = False

    def _alloc_buffers(self, sizes: List[int]) -> None:
        B, dev = self._rows, self.device
        self._X = [torch.empty(B, fi, dtype=torch.bfloat16, device=dev) for (_, fi) in self.dims]
        self._Y = torch.empty(B, self.dims[-1][0], dtype=torch.bfloat16, device=dev)
        self._dZ = [torch.empty(B, fo, dtype=torch.bfloat16, device=dev) for (fo, _) in self.dims]
        if self.boundary is not None:
            if not self.is_first:
                self._X[0] = self.boundary.in_region("act", B, (self.dims[0][1],))
            if not self.is_last:
                self._dZ[-1] = self.boundary.in_region("grad", B, (self.dims[-1][0],))
        self._fp8_bwd_ok = [False] * self.n_layers
        self._XqT = self._dZqT = None
        if self.fp8_backward and getattr(self, "_dp_w", None) is None and all(n % 128 == 0 for n in sizes):
            from ..ops.fp8 import MXFP8Transposed
            self._fp8_bwd_ok = [fo % 128 == 0 and fi % 128 == 0 for (fo, fi) in self.dims]
            self._XqT = [MXFP8Transposed(B, fi, dev) if ok else None
                         for ok, (_, fi) in zip(self._fp8_bwd_ok, self.dims)]
            self._dZqT = [MXFP8Transposed(B, fo, dev) if ok else None
                          for ok, (fo, _) in zip(self._fp8_bwd_ok, self.dims)]
        self._Xq_slots, self._dZq_slots = {}, {}

    def _slice(self, buf: torch.Tensor, mb: int) -> torch.Tensor:
        return buf[self._offs[mb]:self._offs[mb + 1]]

    def input_buffer(self, mb: int) -> Optional[torch.Tensor]:
        """Where the activation of micro-batch ``mb`` should be received (zero-copy)."""
        return self._slice(self._X[0], mb)

    def grad_buffer(self, mb: int) -> Optional[torch.Tensor]:
        """Where the gradient of micro-batch ``mb`` from the next stage should be received."""
        return self._slice(self._dZ[-1], mb)

    # ---- ops ----
    def _requantize_weights(self) -> None:
        """fp8 copies of the (just updated) bf16 weights, written IN PLACE after the first call so
        that nothing is allocated per step and captured graphs keep reading the same memory."""
        from ..ops.fp8 import quantize_mxfp8, quantize_mxfp8_dual
        first = getattr(self, "_wq", None) is None
        if first:
            self._wq, self._wqT = [None] * self.n_layers, [None] * self.n_layers
        for l, (w, (fo, fi)) in enumerate(zip(self.w, self.dims)):
            if self.fp8_backward and fo % 128 == 0 and fi % 128 == 0:
                # W [fo, fi] scaled along fi (forward) and W^T [fi, fo] scaled along fo (dgrad)
                self._wq[l], self._wqT[l] = quantize_mxfp8_dual(
                    w, want_col=True, row_out=self._wq[l], col_out=self._wqT[l])
            elif fi % 128 == 0:
                self._wq[l] = quantize_mxfp8(w, out=self._wq[l])

    def _q_slot(self, table: dict, l: int, mb: int, rows: int, cols: int):
        """Persistent row-wise fp8 slot of (layer, micro-batch) for producer-side quantisation."""
        from ..ops.fp8 import MXFP8Tensor, sf_bytes
        t = table.get((l, mb))
        if t is None or (t.rows, t.cols) != (rows, cols):
            t = table[(l, mb)] = MXFP8Tensor(
                torch.empty(rows, cols, dtype=torch.uint8, device=self.device),
                torch.zeros(sf_bytes(rows, cols), dtype=torch.uint8, device=self.device), rows, cols)
        return t

    def _layer_fwd(self, l: int, x: torch.Tensor, y: torch.Tensor, mb: Optional[int] = None) -> None:
        relu = not (self.last_layer_is_logits and l == self.n_layers - 1)
        if self.fp8_forward and x.shape[1] % 128 == 0:
            from ..ops.fp8 import gemm_mxfp8, quantize_mxfp8, quantize_mxfp8_dual
            if getattr(self, "_wq", None) is None:
                self._requantize_weights()
            if mb is not None and self._fp8_bwd_ok[l]:
                if (l, mb) in self._xq_ready:
                    xq = self._Xq_slots[(l, mb)]         # written by the GEMM that produced x
                else:
                    # training: the same pass also writes x^T (scaled along the rows) for the wgrad GEMM
                    xq, _ = quantize_mxfp8_dual(x, col_into=self._XqT[l], row_offset=self._offs[mb])
            else:
                xq = quantize_mxfp8(x)
            nxt = l + 1
            if mb is not None and self._producer_quant and nxt < self.n_layers and self._fp8_bwd_ok[nxt]:
                # y is the next layer's input: emit its row-wise and transposed fp8 copies here
                gemm_mxfp8(xq, self._wq[l], y, bias=self.b_master[l], relu=relu,
                           q_row_out=self._q_slot(self._Xq_slots, nxt, mb, y.shape[0], y.shape[1]),
                           q_col_into=self._XqT[nxt], q_row_offset=self._offs[mb])
                self._xq_ready.add((nxt, mb))
            else:
                gemm_mxfp8(xq, self._wq[l], y, bias=self.b_master[l], relu=relu)
            return
        if self.fp8_forward and not self._fp8_fallback_logged:
            # say it once: a benchmark line must be able to tell which kernels ran
            import sys
            self._fp8_fallback_logged = True
            print(f"[mlp] stage {self.stage_idx}: layer {l} has in_features {x.shape[1]} (not a multiple of 128): "
                  "bf16 kernels for it instead of MX-fp8", file=sys.stderr, flush=True)
        _gemm(x, self.w[l], y, bias=self.b_master[l], relu=relu)

    def forward(self, mb, x, targets):
        x0 = self._slice(self._X[0], mb)
        if x.data_ptr() != x0.data_ptr():
            x0.copy_(x)                      # first stage: the batch slice (cast to bf16 if needed)
        x = x0
        for l in range(self.n_layers):
            if l + 1 < self.n_layers:
                y = self._slice(self._X[l + 1], mb)
            elif self.boundary is not None and not self.is_last:
                y = self.boundary.out_slot("act", mb, (self.dims[-1][0],))   # the next stage's X[0] rows
            else:
                y = self._slice(self._Y, mb)
            self._layer_fwd(l, x, y, mb)
            x = y
        if self.is_last:
            loss_ops.softmax_ce(x, targets, grad_scale=self._scale, want_grad=True,
                                loss_sum=self.loss_sum, dlogits=self._slice(self._dZ[-1], mb))
            return None
        return x

    def backward_input(self, mb, dy):
        dz = self._slice(self._dZ[-1], mb)
        if not self.is_last and dy.data_ptr() != dz.data_ptr():
            dz.copy_(dy)
        for l in range(self.n_layers - 1, -1, -1):
            no_dgrad = l == 0 and self.is_first
            dzq = None
            if self._fp8_bwd_ok[l]:
                from ..ops.fp8 import gemm_mxfp8, quantize_mxfp8_dual
                if (l, mb) in self._dzq_ready:
                    dzq = self._dZq_slots.get((l, mb))   # written by the dgrad GEMM of the layer above
                else:
                    # one pass over dz_l: row-wise copy for this layer's dgrad, dz_l^T for its wgrad
                    dzq, _ = quantize_mxfp8_dual(dz, want_row=not no_dgrad, col_into=self._dZqT[l],
                                                 row_offset=self._offs[mb])
            if no_dgrad:
                return None
            # dz_{l-1} = (dz_l . W_l) * (x_l > 0): x_l is the post-ReLU output of the layer below
            # (or of the previous stage's last layer), so the mask is its ReLU derivative.
            x_l = self._slice(self._X[l], mb)
            if l > 0:
                prev = self._slice(self._dZ[l - 1], mb)
            elif self.boundary is not None:
                prev = self.boundary.out_slot("grad", mb, (self.dims[0][1],))   # the previous stage's dZ[-1] rows
            else:
                prev = torch.empty(dz.shape[0], self.dims[0][1], dtype=torch.bfloat16, device=self.device)
            if dzq is not None:
                below = l - 1
                if self._producer_quant and below >= 0 and self._fp8_bwd_ok[below]:
                    # prev = dz of the layer below: emit dz^T for its wgrad and, unless that layer has
                    # no dgrad of its own (first layer of the model), the row-wise copy for its dgrad
                    need_row = not (below == 0 and self.is_first)
                    gemm_mxfp8(dzq, self._wqT[l], prev, aux=x_l,
                               q_row_out=(self._q_slot(self._dZq_slots, below, mb, prev.shape[0], prev.shape[1])
                                          if need_row else None),
                               q_col_into=self._dZqT[below], q_row_offset=self._offs[mb])
                    self._dzq_ready.add((below, mb))
                else:
                    gemm_mxfp8(dzq, self._wqT[l], prev, aux=x_l)  # K = fan_out, B operand = W^T [fan_in, fan_out]
            else:
                _gemm(dz, self.w[l], prev, b_mn=1, aux=x_l)
            dz = prev
        return dz

    # ---- data parallel: wgrad GEMM fused with reduce-scatter + sharded SGD + weight push ----
    def enable_fused_dp(self, groups, topo, rank) -> None:
        """Switch this stage's DP reduction to the fused wgrad -> reduce-scatter -> sharded SGD -> weight
        push kernels.  The decision is GLOBAL: building the symmetric buffers is collective over the
        whole job, so every rank votes and the fused path is taken only if every stage of every
        replica can take it; otherwise all of them keep the NCCL all-reduce -- and say so."""
        import sys
        import torch.distributed as dist
        from ..ops.optim import FusedSGD
        from ..parallel.dp_ops import WgradReduceScatterSGD
        from ..parallel.symm import FusedDPSGD
        dp_ranks = topo.dp_group_ranks(rank)
        dp = len(dp_ranks)
        why = None
        if self.device.type != "cuda":
            why = "not on CUDA"
        elif not isinstance(self.optimizer, FusedSGD):
            why = "optimizer is not SGD"
        elif self.w_group != 0 or self.pp > 1:
            # the fused kernels reduce ONE whole-step wgrad GEMM per layer; pipelines cut the step's
            # weight gradients into groups so that W ops can fill bubbles
            why = f"pipeline stages run {self.w_group or 'several'} micro-batches per wgrad group (needs one per step)"
        elif any(fo % (32 * dp) or fi % 8 for (fo, fi) in self.dims):
            why = f"a layer dimension is not a multiple of 32*dp = {32 * dp} (out) / 8 (in): {self.dims}"
        vote = torch.tensor([0 if why is None else 1], dtype=torch.int32)
        dist.all_reduce(vote, group=groups.ctrl)
        if int(vote.item()) != 0:
            if rank == 0 or why is not None:
                print(f"[mlp] rank {rank}: fused DP path OFF, NCCL all-reduce + local SGD instead"
                      + (f": {why}" if why else " (another stage cannot take it)"), file=sys.stderr, flush=True)
            return
        opt = self.optimizer
        self._dp_w = []
        for l, (fo, fi) in enumerate(self.dims):
            op = WgradReduceScatterSGD(dp_ranks, rank, groups.ctrl, fo, fi, self.device,
                                       self.w_master[l], opt.momentum_buffers[l], self.w[l])
            self.w[l] = op.weight          # compute copy now lives in symmetric memory
            self._dp_w.append(op)
        nl = self.n_layers
        self._dp_b = FusedDPSGD(self.b_master, lambda: self.db, opt.momentum_buffers[nl:], opt.lr,
                                opt.momentum, opt.weight_decay, dp_ranks, rank, groups.ctrl, self.device,
                                timeout_s=getattr(self, "timeout_s", 300.0))
        self._dp_ranks, self._dp_groups = dp_ranks, groups
        if rank == 0:
            print("[mlp] fused DP path ON: wgrad GEMM -> reduce-scatter -> sharded SGD -> weight push",
                  file=sys.stderr, flush=True)

    def _flush_w_fused_dp(self, lo: int, hi: int) -> None:
        opt = self.optimizer
        pars = []
        for l in range(self.n_layers):
            dz, x = self._dZ[l][lo:hi], self._X[l][lo:hi]
            pars.append(self._dp_w[l].wgrad(dz, x, 1.0 / self.dp_size))
            _colsum(dz, self.db[l], accumulate=False)
        for l in range(self.n_layers):
            self._dp_w[l].finalize_sgd(pars[l], opt.lr, opt.momentum, opt.weight_decay)
        self._dp_b.step(grad_scale=1.0 / self.dp_size)
        for l in range(self.n_layers):
            self._dp_w[l].wait_refreshed()
        self._dp_step_done = True

    def _flush_w(self) -> None:
        if not self._pending_w:
            return
        lo, hi = self._offs[self._pending_w[0]], self._offs[self._pending_w[-1] + 1]
        whole_step = (not self._w_done_any) and lo == 0 and hi == self._offs[-1]
        if getattr(self, "_dp_w", None) is not None and whole_step:
            self._flush_w_fused_dp(lo, hi)
            self._w_done_any = True
            self._pending_w = []
            return
        acc = self._w_done_any
        for l in range(self.n_layers):
            dz, x = self._dZ[l][lo:hi], self._X[l][lo:hi]
            # dW_l (+)= dz^T . x   (A = dz stored [K=rows, M=out], B = x stored [K=rows, N=in])
            if self._fp8_bwd_ok[l]:
                from ..ops.fp8 import gemm_mxfp8
                gemm_mxfp8(self._dZqT[l].rows_slice(lo, hi), self._XqT[l].rows_slice(lo, hi), self.dw[l],
                           accumulate=acc)
            else:
                _gemm(dz, x, self.dw[l], a_mn=1, b_mn=1, accumulate=acc)
            _colsum(dz, self.db[l], accumulate=acc)
        self._w_done_any = True
        self._pending_w = []

    def flush_weight_grads(self) -> None:
        """Run the weight-gradient GEMMs of every micro-batch queued by ``backward_weight``."""
        self._flush_w()

    def backward_weight(self, mb) -> None:
        if self._pending_w and mb != self._pending_w[-1] + 1:
            self._flush_w()                  # only consecutive micro-batches share one GEMM
        self._pending_w.append(mb)
        if self.w_group and len(self._pending_w) >= self.w_group:
            self._flush_w()

    def end_step(self) -> None:
        self._dp_step_done = False
        self._flush_w()
        if self._dp_step_done:
            self.optimizer.step_count += 1     # reduction + update already done inside the kernels
            if self.fp8_forward:
                self._requantize_weights()
            return
        if getattr(self, "_dp_w", None) is not None:
            raise RuntimeError("fused DP needs one wgrad group per step (w_group = 0)")
        grad_scale = 1.0
        if self.dp_group is not None and self.dp_size > 1:
            import torch.distributed as dist
            for g in self.dw + self.db:
                dist.all_reduce(g, group=self.dp_group)
            grad_scale = 1.0 / self.dp_size
        # the first wgrad of a step overwrites dW, so nothing has to be zeroed here
        self.optimizer.step(grad_scale=grad_scale, zero_grad=False)
        if self.fp8_forward:
            self._requantize_weights()

    @torch.no_grad()
    def eval_forward(self, x, targets):
        if x.dtype != torch.bfloat16:
            x = x.to(torch.bfloat16)
        for l, (fo, _) in enumerate(self.dims):
            y = torch.empty(x.shape[0], fo, dtype=torch.bfloat16, device=self.device)
            self._layer_fwd(l, x, y)
            x = y
        y = x
        if self.is_last:
            loss_ops.softmax_ce(y, targets, want_grad=False, want_correct=True,
                                loss_sum=self.loss_sum, correct=self.correct)
            return None
        return y

    # ---- checkpoint: nn.Module-style names so a plain nn.Sequential-of-Linear can load it ----
    def _gather_dp_shards(self) -> None:
        """Fused DP keeps master rows AND their momentum current only on the owning replica: collect
        the shards on every replica (collective over the DP group; rare, NCCL is fine)."""
        if getattr(self, "_dp_w", None) is None:
            return
        import torch.distributed as dist
        for l, op in enumerate(self._dp_w):
            for full in (self.w_master[l], self.optimizer.momentum_buffers[l]):
                shard = full[op.idx * op.rows:(op.idx + 1) * op.rows].contiguous()
                dist.all_gather_into_tensor(full, shard, group=self.dp_group)

    def optimizer_state(self):
        """Optimizer state for the checkpoint (every DP replica must call it: see ``_gather_dp_shards``)."""
        self._gather_dp_shards()
        return self.optimizer.state_dict()

    def state_dict(self):
        sd = {}
        self._gather_dp_shards()
        for l in range(self.n_layers):
            sd[f"layers.{l}.weight"] = self.w_master[l].detach().cpu().clone()
            sd[f"layers.{l}.bias"] = self.b_master[l].detach().cpu().clone()
        return sd

    def load_state_dict(self, sd):
        for l in range(self.n_layers):
            self.w_master[l].copy_(sd[f"layers.{l}.weight"])
            self.b_master[l].copy_(sd[f"layers.{l}.bias"])
            self.w[l].copy_(self.w_master[l])
        if getattr(self, "_wq", None) is not None:
            self._requantize_weights()       # fp8 copies follow the loaded weights (in place)

    def flops_per_sample(self) -> int:
        """Model FLOPs per sample for fwd + dgrad + wgrad of this stage (2*M*N*K each)."""
        total = 0
        for l, (fo, fi) in enumerate(self.dims):
            n_gemm = 3 if not (l == 0 and self.is_first) else 2
            total += n_gemm * 2 * fo * fi
        return total
