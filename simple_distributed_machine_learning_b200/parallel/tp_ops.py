"""Tensor-parallel fused ops: all-gather -> GEMM and GEMM -> reduce-scatter
(``csrc/gemm_tcgen05.cu``, CTA-pair tcgen05 kernel with communication hooks).

``AllGatherGemm`` owns one row-stacked operand buffer ``[M, K]`` in symmetric memory per
use site; a rank writes its own rows ``[rank*M/tp, (rank+1)*M/tp)`` (e.g. by receiving its
shard from the previous pipeline stage straight into them) and calls ``run``: the kernel
pulls the other ranks' rows over NVLink while it already multiplies the local ones.  The
gathered operand stays in the buffer for the weight-gradient GEMM.

``GemmReduceScatter`` owns the bf16 landing slots ``[tp][M/tp, N]`` (two sets, alternating
by call parity): every rank's kernel TMA-stores the rows of its partial result into the slot
it owns on the GPU that owns those rows, ``finalize`` waits for all contributions, sums
them in fp32 and applies bias / ReLU / mask.

No counterpart in the reference (it never shards a layer; /root/reference/simple_distributed.py:26-83
splits the model only BETWEEN conv2 and fc1); these are the tensor-parallel forms of its
Linear+bias+ReLU layers (ref:63-64, :75-77).
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

from ..ops import _lib
from .symm import SymmetricBuffer


def _ptr_array(ptrs: List[int]):
    return (ctypes.c_void_p * len(ptrs))(*ptrs)


class AllGatherGemm:
    """``slots`` independent [M, K] operand regions (one per micro-batch in flight), stacked
    slot-major so that consecutive slots form one contiguous [slots*M, K] matrix -- which is
    what the grouped weight-gradient GEMM reads."""

    MAX_CHUNKS = 64

    def __init__(self, tp_ranks: List[int], my_rank: int, ctrl_group, M: int, K: int,
                 device: torch.device, chunk_rows: int = 256, slots: int = 1):
        self.tp = len(tp_ranks)
        self.idx = tp_ranks.index(my_rank)
        self.M, self.K, self.slots = M, K, slots
        assert M % (256 * self.tp) == 0, "rows must split into 256-row blocks per rank"
        self.rows_per_rank = M // self.tp
        self.chunk_rows = chunk_rows
        assert self.rows_per_rank % chunk_rows == 0 and self.rows_per_rank // chunk_rows <= self.MAX_CHUNKS
        self.slot_bytes = M * K * 2
        self.flags_off = ((slots * self.slot_bytes + 255) // 256) * 256
        self.count_off = self.flags_off + 256 * slots
        self.count_bytes = 4 * self.tp * self.MAX_CHUNKS
        total = self.count_off + self.count_bytes * slots
        self.buf = SymmetricBuffer(tp_ranks, my_rank, total, ctrl_group, device)
        self.all = self.buf.local_view(torch.bfloat16, 0, slots * M * K).view(slots * M, K)
        self.full = self.all[:M]
        self.epochs = [0] * slots
        self._src = [_ptr_array([self.buf.ptr(r, s * self.slot_bytes) for r in range(self.tp)])
                     for s in range(slots)]
        self._flags = [_ptr_array([self.buf.ptr(r, self.flags_off + 256 * s) for r in range(self.tp)])
                       for s in range(slots)]
        self._count = [ctypes.c_void_p(self.buf.ptr(self.idx, self.count_off + self.count_bytes * s))
                       for s in range(slots)]

    def slot_rows(self, slot: int = 0) -> torch.Tensor:
        """The full [M, K] operand region of ``slot`` (valid after :meth:`run`)."""
        return self.all[slot * self.M:(slot + 1) * self.M]

    def local_rows(self, slot: int = 0) -> torch.Tensor:
        """This rank's shard of the operand (write the rows to be gathered here)."""
        lo = slot * self.M + self.idx * self.rows_per_rank
        return self.all[lo:lo + self.rows_per_rank]

    def run(self, W: torch.Tensor, C: torch.Tensor, *, b_mn: int = 0, bias=None, aux=None,
            relu: bool = False, alpha: float = 1.0, slot: int = 0) -> torch.Tensor:
        """C[M, N] = epilogue(gathered A . op(W)^T).  ``b_mn=0``: W is [N, K]; ``b_mn=1``: [K, N]."""
        N = C.shape[1]
        assert C.shape[0] == self.M and C.stride(1) == 1
        self.epochs[slot] += 1
        A = self.slot_rows(slot)
        lib = _lib.cuda_lib()
        rc = lib.sdml_gemm_bf16_comm(
            A.data_ptr(), W.data_ptr(), C.data_ptr(), self.M, N, self.K, A.stride(0),
            W.stride(0), C.stride(0), int(b_mn), bias.data_ptr() if bias is not None else None,
            aux.data_ptr() if aux is not None else None, aux.stride(0) if aux is not None else 0,
            int(relu), int(C.dtype == torch.float32), float(alpha),
            self.tp, self.idx, self._src[slot], self._flags[slot], self._count[slot], self.MAX_CHUNKS,
            self.chunk_rows, ctypes.c_uint(self.epochs[slot]), 1, 0, None, 0, None,
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_gemm_bf16_comm(all-gather)")
        return C


class GemmReduceScatter:
    """Landing buffers: for each call parity, ``tp`` slots of bf16 ``[M/tp, N]`` in this rank's
    symmetric memory; slot ``s`` is written by rank ``s``'s kernel (plain TMA stores, remote
    ones over NVLink), ``finalize`` sums the slots."""

    def __init__(self, tp_ranks: List[int], my_rank: int, ctrl_group, M: int, N: int,
                 device: torch.device):
        self.tp = len(tp_ranks)
        self.idx = tp_ranks.index(my_rank)
        self.M, self.N = M, N
        assert M % (32 * self.tp) == 0 and N % 8 == 0
        self.rows = M // self.tp
        self.slot_elems = self.rows * N
        self.slot_bytes = ((self.slot_elems * 2 + 255) // 256) * 256
        self.par_bytes = self.slot_bytes * self.tp
        self.done_off = 2 * self.par_bytes
        self.buf = SymmetricBuffer(tp_ranks, my_rank, self.done_off + 512, ctrl_group, device)
        self.calls = 0
        self.grid = _lib.cuda_lib().sdml_gemm_comm_grid()
        # destination for owner r = r's parity region, slot `self.idx`
        self._dst = [_ptr_array([self.buf.ptr(r, par * self.par_bytes + self.idx * self.slot_bytes)
                                 for r in range(self.tp)]) for par in (0, 1)]
        # one completion counter per parity so that a fast peer's next call cannot be mistaken
        self._done = [_ptr_array([self.buf.ptr(r, self.done_off + 256 * par) for r in range(self.tp)])
                      for par in (0, 1)]
        self._uses = [0, 0]

    def run(self, A: torch.Tensor, W: torch.Tensor, *, b_mn: int = 0, alpha: float = 1.0) -> int:
        """Partial = A[M, K] . op(W)^T, rows stored into their owners' landing slots.  Returns
        the parity token to pass to :meth:`finalize`."""
        M, K = A.shape
        assert M == self.M and A.stride(1) == 1
        par = self.calls & 1
        self.calls += 1
        self._uses[par] += 1
        lib = _lib.cuda_lib()
        rc = lib.sdml_gemm_bf16_comm(
            A.data_ptr(), W.data_ptr(), None, M, self.N, K, A.stride(0), W.stride(0), self.N, int(b_mn),
            None, None, 0, 0, 0, float(alpha),
            1, 0, None, None, None, 0, 256, ctypes.c_uint(0),
            self.tp, self.idx, self._dst[par], self.N, self._done[par],
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_gemm_bf16_comm(reduce-scatter)")
        return par

    def finalize(self, par: int, out: torch.Tensor, *, bias=None, aux=None, relu: bool = False) -> torch.Tensor:
        """out[M/tp, N] (bf16) = mask(relu(sum of the landed partials + bias))."""
        assert out.shape == (self.rows, self.N) and out.dtype == torch.bfloat16
        expected = self._uses[par] * self.tp * 4 * self.grid
        lib = _lib.cuda_lib()
        rc = lib.sdml_rs_finalize(
            self.buf.ptr(self.idx, par * self.par_bytes), self.tp, self.slot_bytes // 2, self.N,
            self.rows, self.N, bias.data_ptr() if bias is not None else None,
            aux.data_ptr() if aux is not None else None, aux.stride(0) if aux is not None else 0,
            int(relu), out.data_ptr(), out.stride(0),
            self.buf.ptr(self.idx, self.done_off + 256 * par), ctypes.c_uint(expected),
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_rs_finalize")
        return out


class Fp8GemmReduceScatter(GemmReduceScatter):
    """GEMM -> reduce-scatter with MX-fp8 operands (``sdml_gemm_mxfp8_2cta_rs``): the block-scaled
    CTA-pair kernel TMA-stores the rows of its bf16 partial result into the owners' landing slots; the
    landing buffers, the completion counters and :meth:`finalize` are the bf16 class's.  Replaces the
    fp32 partial + ``reduce_scatter_tensor`` of the unfused fp8 path (a [rows, N] fp32 tensor written,
    read and sent per call)."""

    def run(self, Aq, Wq, *, alpha: float = 1.0) -> int:
        """``Aq``: MXFP8Tensor [M, K] (scales along K); ``Wq``: MXFP8Tensor [N, K].  Returns the parity."""
        assert Aq.rows == self.M and Wq.rows == self.N and Aq.cols == Wq.cols
        par = self.calls & 1
        self.calls += 1
        self._uses[par] += 1
        rc = _lib.cuda_lib().sdml_gemm_mxfp8_2cta_rs(
            Aq.data.data_ptr(), Aq.scales.data_ptr(), Wq.data.data_ptr(), Wq.scales.data_ptr(),
            self.M, self.N, Aq.cols, Aq.data.stride(0), Wq.data.stride(0), float(alpha),
            self.tp, self.idx, self._dst[par], self.N, self._done[par],
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_gemm_mxfp8_2cta_rs")
        return par


class Fp8AllGather:
    """All-gather of an MX-fp8 operand by PUSH: every tensor-parallel rank quantises only ITS row
    shard (row-wise copy for the forward / dgrad GEMM, transposed copy for the weight-gradient GEMM),
    straight into its own rows of the operand buffers, and copies those fp8 bytes + scale atoms into the
    same place of every peer's buffers over NVLink on the copy engine (side stream); a flag per (slot,
    source rank) tells the consumer's stream that the peer's rows have landed.  Compared with gathering
    bf16 rows through NCCL and quantising the whole operand on every rank this moves the same bytes,
    quantises 1/tp of them and needs no collective call.

    Buffers (symmetric, per rank): ``slots`` operands [M, K] row-wise (``rows_q(slot)``) and ONE
    step-wide transposed operand [K, slots * M] (``transposed``, an ``MXFP8Transposed`` view: the wgrad
    GEMM K-slices it exactly like the unfused path's buffer).  Slot reuse across steps is guarded by an
    acknowledgement per rank and step (``end_step``).  Host-side epochs: the tensor-parallel stages run
    on the eager op-list executor."""

    def __init__(self, tp_ranks: List[int], my_rank: int, ctrl_group, M: int, K: int, device: torch.device,
                 slots: int = 1, timeout_s: float = 300.0):
        from ..ops.fp8 import MXFP8Tensor, MXFP8Transposed, sf_bytes
        from .boundary import device_bytes
        self.tp = len(tp_ranks)
        self.idx = tp_ranks.index(my_rank)
        self.M, self.K, self.slots, self.device = M, K, slots, device
        assert M % (128 * self.tp) == 0 and K % 128 == 0
        self.rs = M // self.tp
        self.groups = M // 128                      # row groups of one row-wise operand
        self.kb = K // 128
        self.q_bytes = M * K
        self.sf_slot = sf_bytes(M, K)               # = kb * groups * 512
        B = slots * M
        self.B = B
        self.off_q = 0
        self.off_sf = self.off_q + slots * self.q_bytes
        self.off_qt = ((self.off_sf + slots * self.sf_slot + 255) // 256) * 256
        self.off_sft = self.off_qt + K * B
        self.sft_bytes = sf_bytes(K, B)             # = (B / 128) * kb * 512
        self.off_flags = ((self.off_sft + self.sft_bytes + 255) // 256) * 256      # [slots][tp] uint32
        self.off_ack = self.off_flags + 4 * slots * self.tp                          # [tp] uint32
        total = ((self.off_ack + 4 * self.tp + 255) // 256) * 256
        self.buf = SymmetricBuffer(tp_ranks, my_rank, total, ctrl_group, device)
        self.lib = _lib.cuda_lib()
        lb = self.buf.local_bytes
        self._q = [lb[self.off_q + s * self.q_bytes: self.off_q + (s + 1) * self.q_bytes].view(M, K) for s in range(slots)]
        self._sf = [lb[self.off_sf + s * self.sf_slot: self.off_sf + (s + 1) * self.sf_slot] for s in range(slots)]
        self._mx = [MXFP8Tensor(self._q[s], self._sf[s], M, K) for s in range(slots)]
        t = MXFP8Transposed.__new__(MXFP8Transposed)
        t.total_rows, t.cols = B, K
        t.data = lb[self.off_qt: self.off_qt + K * B].view(K, B)
        t.groups = self.kb
        t.scales = lb[self.off_sft: self.off_sft + self.sft_bytes]
        self.transposed = t
        self._zero = torch.zeros(1, dtype=torch.int32, device=device)        # "counter" of the flag kernels
        self._err_dev = torch.zeros(4, dtype=torch.int32, device=device)
        self._err_host = torch.zeros(4, dtype=torch.int32).pin_memory()
        self.timeout_ns = int(max(0.0, timeout_s) * 1e9)
        self._uses = [0] * slots
        self._step = 0
        self._acked = False
        self._side = torch.cuda.Stream(device=device)
        self._ev = torch.cuda.Event()
        self._peers = [r for r in range(self.tp) if r != self.idx]
        del device_bytes

    def rows_q(self, slot: int):
        """The full row-wise operand of ``slot`` (valid on the current stream after :meth:`gather`)."""
        return self._mx[slot]

    def _p(self, r: int, off: int) -> int:
        return self.buf.ptr(r, off)

    def gather(self, slot: int, shard: torch.Tensor):
        """``shard``: this rank's bf16 rows [M / tp, K] of the operand.  Returns the full MXFP8Tensor."""
        assert shard.shape == (self.rs, self.K) and shard.dtype == torch.bfloat16 and shard.stride(1) == 1
        lib, me, rs, K, M, B = self.lib, self.idx, self.rs, self.K, self.M, self.B
        cur = torch.cuda.current_stream(self.device)
        r0 = me * rs                                   # first row of my shard inside the slot's operand
        c0 = slot * M + r0                             # first column of my shard inside the transposed operand
        g0 = r0 // 128
        # 1. quantise my shard in place (both copies) on the compute stream
        rc = lib.sdml_quant_mxfp8_dual_at(
            shard.data_ptr(), shard.stride(0), rs, K,
            self._p(me, self.off_q + slot * self.q_bytes + r0 * K), K,
            self._p(me, self.off_sf + slot * self.sf_slot + g0 * 512), self.groups,
            self._p(me, self.off_qt + c0), B,
            self._p(me, self.off_sft + (c0 // 128) * self.kb * 512), self.kb, cur.cuda_stream)
        _lib.check(rc, "sdml_quant_mxfp8_dual_at")
        self._uses[slot] += 1
        use = self._uses[slot]
        # 2. push my fp8 rows + scale atoms into every peer's buffers (copy engine, side stream)
        self._ev.record(cur)
        side = self._side
        side.wait_event(self._ev)
        ss = side.cuda_stream
        for r in self._peers:
            if not self._acked:
                # first push of a step: the peer must have finished reading last step's operands
                _lib.check(lib.sdml_bnd_wait(self._p(me, self.off_ack + 4 * r), self._zero.data_ptr(), self._step,
                                             self.timeout_ns, self._err_dev.data_ptr(), self._err_host.data_ptr(),
                                             16, ss), "sdml_bnd_wait")
            copies = (
                # row-wise payload: rs contiguous rows
                (self.off_q + slot * self.q_bytes + r0 * K, rs * K, rs * K, rs * K, 1),
                # row-wise scale atoms: for every k-block a run of rs / 128 atoms
                (self.off_sf + slot * self.sf_slot + g0 * 512, self.groups * 512, self.groups * 512,
                 (rs // 128) * 512, self.kb),
                # transposed payload: a column range of every one of the K rows
                (self.off_qt + c0, B, B, rs, K),
                # transposed scale atoms: the row blocks of my columns, contiguous
                (self.off_sft + (c0 // 128) * self.kb * 512, (rs // 128) * self.kb * 512,
                 (rs // 128) * self.kb * 512, (rs // 128) * self.kb * 512, 1),
            )
            for off, dpitch, spitch, width, height in copies:
                _lib.check(lib.sdml_memcpy2d_async(self._p(r, off), dpitch, self._p(me, off), spitch, width, height, ss),
                           "sdml_memcpy2d_async")
            flag = self._p(r, self.off_flags + 4 * (slot * self.tp + me))
            _lib.check(lib.sdml_bnd_signal(flag, self._zero.data_ptr(), use, 0, ss), "sdml_bnd_signal")
        self._acked = True
        # 3. the consumer (current stream) waits for every peer's rows of this use
        for r in self._peers:
            _lib.check(lib.sdml_bnd_wait(self._p(me, self.off_flags + 4 * (slot * self.tp + r)), self._zero.data_ptr(),
                                         use, self.timeout_ns, self._err_dev.data_ptr(), self._err_host.data_ptr(),
                                         17, cur.cuda_stream), "sdml_bnd_wait")
        return self._mx[slot]

    def end_step(self) -> None:
        """After the last reader of this step's operands has been enqueued on the current stream:
        tell every peer that it may overwrite this rank's slots, and join the push stream."""
        cur = torch.cuda.current_stream(self.device)
        self._step += 1
        for r in self._peers:
            _lib.check(self.lib.sdml_bnd_signal(self._p(r, self.off_ack + 4 * self.idx), self._zero.data_ptr(),
                                                self._step, 0, cur.cuda_stream), "sdml_bnd_signal")
        cur.wait_stream(self._side)          # my pushes read my buffers: keep them ordered before next step's writes
        self._acked = False

    def check(self) -> None:
        if int(self._err_host[0]) != 0:
            what = "acknowledgement" if int(self._err_host[0]) == 16 else "fp8 rows"
            raise RuntimeError(f"tensor-parallel fp8 all-gather: gave up waiting for a peer's {what} "
                               f"(flag {int(self._err_host[1]) & 0xFFFFFFFF}, wanted {int(self._err_host[2]) & 0xFFFFFFFF})")
