"""Data-parallel weight-gradient reduction fused into the wgrad GEMM + sharded optimizer
(``sdml_wgrad_rs`` / ``sdml_rs_finalize_sgd`` in ``csrc/gemm_tcgen05.cu``).

For one weight ``[out, in]`` replicated over ``dp`` GPUs:

* every replica's wgrad GEMM stores its (scaled, bf16) tiles into the landing slot it owns
  on the replica that owns those rows -- a reduce-scatter over NVSwitch hidden behind the MMA
  main loop;
* the owner sums the ``dp`` slots in fp32, applies SGD-with-momentum to its rows of the fp32
  master weight (master + momentum are thereby sharded ZeRO-1 style) and writes the new bf16
  rows into every replica's compute copy with P2P stores -- the all-gather half;
* a counter tells every replica when all owners have refreshed it.

The reference has no data parallelism (one loader on rank 0, /root/reference/simple_distributed.py:94;
extra ranks idle, ref:177-184); this replaces what a DDP all-reduce + ``DistributedOptimizer.step``
(ref:100-104, :113) would do for replicated stages.
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

from ..ops import _lib
from .symm import SymmetricBuffer


def _ptr_array(ptrs: List[int]):
    return (ctypes.c_void_p * len(ptrs))(*ptrs)


class WgradReduceScatterSGD:
    def __init__(self, dp_ranks: List[int], my_rank: int, ctrl_group, out_f: int, in_f: int,
                 device: torch.device, master: torch.Tensor, momentum_buf: torch.Tensor,
                 init_bf16: torch.Tensor):
        self.dp = len(dp_ranks)
        self.idx = dp_ranks.index(my_rank)
        self.out_f, self.in_f = out_f, in_f
        assert out_f % (32 * self.dp) == 0 and in_f % 8 == 0
        self.rows = out_f // self.dp
        self.slot_bytes = self.rows * in_f * 2
        self.par_bytes = self.slot_bytes * self.dp
        self.w_off = 2 * self.par_bytes
        self.w_bytes = out_f * in_f * 2
        self.ctr_off = ((self.w_off + self.w_bytes + 255) // 256) * 256
        # counters: done[par 0], done[par 1], ready, exit  (256 B apart)
        self.buf = SymmetricBuffer(dp_ranks, my_rank, self.ctr_off + 4 * 256, ctrl_group, device)
        self.weight = self.buf.local_view(torch.bfloat16, self.w_off, out_f * in_f).view(out_f, in_f)
        self.weight.copy_(init_bf16)
        self.master, self.mom = master, momentum_buf
        self.calls = 0
        self._uses = [0, 0]
        self.grid = _lib.cuda_lib().sdml_gemm_comm_grid()
        self._dst = [_ptr_array([self.buf.ptr(r, par * self.par_bytes + self.idx * self.slot_bytes)
                                 for r in range(self.dp)]) for par in (0, 1)]
        self._done = [_ptr_array([self.buf.ptr(r, self.ctr_off + 256 * par) for r in range(self.dp)])
                      for par in (0, 1)]
        row_off = self.idx * self.rows * in_f * 2
        self._wout = _ptr_array([self.buf.ptr(r, self.w_off + row_off) for r in range(self.dp)])
        self._ready = _ptr_array([self.buf.ptr(r, self.ctr_off + 512) for r in range(self.dp)])
        self.ready_ptr = self.buf.ptr(self.idx, self.ctr_off + 512)
        self.finalized = 0

    def wgrad(self, dz: torch.Tensor, x: torch.Tensor, alpha: float) -> int:
        """Launch dW = alpha * dz^T . x with the reduce-scatter epilogue; returns the parity."""
        par = self.calls & 1
        self.calls += 1
        self._uses[par] += 1
        rc = _lib.cuda_lib().sdml_wgrad_rs(
            dz.data_ptr(), x.data_ptr(), self.out_f, self.in_f, dz.shape[0], dz.stride(0), x.stride(0),
            float(alpha), self.dp, self.idx, self._dst[par], self.in_f, self._done[par],
            torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_wgrad_rs")
        return par

    def finalize_sgd(self, par: int, lr: float, momentum: float, weight_decay: float) -> None:
        expected = self._uses[par] * self.dp * 4 * self.grid
        lo = self.idx * self.rows
        m = self.master[lo:lo + self.rows]
        mo = self.mom[lo:lo + self.rows]
        rc = _lib.cuda_lib().sdml_rs_finalize_sgd(
            self.buf.ptr(self.idx, par * self.par_bytes), self.dp, self.slot_bytes // 2, self.in_f,
            self.rows, self.in_f, m.data_ptr(), mo.data_ptr(), self.master.stride(0),
            float(lr), float(momentum), float(weight_decay), self._wout, self.in_f, self.dp,
            self.buf.ptr(self.idx, self.ctr_off + 256 * par), ctypes.c_uint(expected), self._ready,
            self.buf.ptr(self.idx, self.ctr_off + 768), torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_rs_finalize_sgd")
        self.finalized += 1

    def wait_refreshed(self) -> None:
        """Hold the stream until every owner has pushed its refreshed rows into this replica."""
        rc = _lib.cuda_lib().sdml_wait_counter(self.ready_ptr, ctypes.c_uint(self.finalized * self.dp),
                                               torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_wait_counter")
