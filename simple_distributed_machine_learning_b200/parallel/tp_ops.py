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
