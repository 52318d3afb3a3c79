"""Symmetric (peer-mapped) GPU buffers for the in-kernel collectives (``csrc/symm.cu``).

``SymmetricBuffer(group, nbytes)`` allocates ``nbytes`` on every rank of ``group``,
exchanges CUDA IPC handles through a CPU (gloo) group and maps every peer's buffer, so
kernels can address ``buf.ptrs[r]`` (rank ``r``'s buffer) with plain loads / stores /
reductions that travel over NVLink through the NVSwitch.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib


class _CudaArray:
    """Minimal ``__cuda_array_interface__`` holder so torch can view raw device memory."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2}


class SymmetricBuffer:
    def __init__(self, ranks: List[int], my_rank: int, nbytes: int, ctrl_group=None,
                 device: Optional[torch.device] = None):
        lib = _lib.cuda_lib()
        self._lib = lib
        self.ranks = list(ranks)
        self.world = len(self.ranks)
        self.index = self.ranks.index(my_rank)
        self.nbytes = int(nbytes)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        local = ctypes.c_void_p()
        _lib.check(lib.sdml_symm_alloc(ctypes.c_longlong(self.nbytes), ctypes.byref(local)), "sdml_symm_alloc")
        self.local_ptr = int(local.value)
        handle = (ctypes.c_ubyte * 64)()
        _lib.check(lib.sdml_symm_export(ctypes.c_void_p(self.local_ptr), handle), "sdml_symm_export")
        mine = bytes(handle)
        self.ptrs: List[int] = [0] * self.world
        self._imported: List[int] = []
        if self.world > 1:
            gathered: List[Optional[tuple]] = [None] * dist.get_world_size(ctrl_group)
            dist.all_gather_object(gathered, (my_rank, mine), group=ctrl_group)
            by_rank = {r: h for (r, h) in gathered}
            for i, r in enumerate(self.ranks):
                if r == my_rank:
                    self.ptrs[i] = self.local_ptr
                else:
                    buf = (ctypes.c_ubyte * 64).from_buffer_copy(by_rank[r])
                    out = ctypes.c_void_p()
                    _lib.check(lib.sdml_symm_import(buf, ctypes.byref(out)), "sdml_symm_import")
                    self.ptrs[i] = int(out.value)
                    self._imported.append(int(out.value))
        else:
            self.ptrs[0] = self.local_ptr
        self._holder = _CudaArray(self.local_ptr, self.nbytes)
        self.local_bytes = torch.as_tensor(self._holder, device=self.device)

    def local_view(self, dtype: torch.dtype, offset_bytes: int, numel: int) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        return self.local_bytes[offset_bytes:offset_bytes + numel * esz].view(dtype)

    def ptr(self, r: int, offset_bytes: int = 0) -> int:
        return self.ptrs[r] + offset_bytes

    def close(self) -> None:
        for p in self._imported:
            self._lib.sdml_symm_unimport(ctypes.c_void_p(p))
        self._imported = []
        if self.local_ptr:
            self._lib.sdml_symm_free(ctypes.c_void_p(self.local_ptr))
            self.local_ptr = 0


class FusedDPSGD:
    """All-reduce over the DP replicas of a stage fused with SGD-with-momentum, one kernel
    per step (``csrc/dp_allreduce_sgd.cu``).  Drop-in for ``FusedSGD.step`` on small stages.

    Default protocol: "ll" -- every replica pushes (value, epoch) pairs into its slot of every
    peer's buffer and polls its own memory (one one-way NVLink trip).  ``SDML_DP_PROTOCOL=staged``
    selects the staging-buffer + flag protocol (two round trips; kept as the reference form)."""

    def __init__(self, params: List[torch.Tensor], grads_fn, momentum_buffers: List[torch.Tensor],
                 lr: float, momentum: float, weight_decay: float, dp_ranks: List[int], my_rank: int,
                 ctrl_group, device: torch.device, timeout_s: float = 300.0):
        self.params, self.grads_fn, self.moms = params, grads_fn, momentum_buffers
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.world = len(dp_ranks)
        self.total = sum(p.numel() for p in params)
        self.protocol = os.environ.get("SDML_DP_PROTOCOL", "ll")
        self.ll_bytes = 2 * len(dp_ranks) * self.total * 8
        self.stage_bytes = ((self.total * 4 + 255) // 256) * 256
        # layout: [stage parity 0][stage parity 1][flags: world x u32 (256 B)][counter (256 B)]
        self.flags_off = 2 * self.stage_bytes
        self.counter_off = self.flags_off + 256
        self.ll_off = self.counter_off + 256
        self.buf = SymmetricBuffer(dp_ranks, my_rank, self.ll_off + self.ll_bytes, ctrl_group, device)
        self.index = self.buf.index
        self.epoch = 0
        # the step counter lives on the device so that a CUDA-graph replay of the step sees the
        # right epoch / staging parity; it is advanced by a stream-ordered add after the kernel
        # [0] completed calls, advanced by the LL kernel itself (its last CTA); [1] CTA exit counter
        self.epoch_dev = torch.zeros(2, dtype=torch.int32, device=device)
        self.timeout_ns = int(max(0.0, timeout_s) * 1e9)
        self.err_host = torch.zeros(4, dtype=torch.int32).pin_memory()   # written by a poll that gave up
        self.stamps = None
        if os.environ.get("SDML_DP_STAMPS", "0") == "1":
            self.stamps = torch.zeros(64 * 4, dtype=torch.int64, device=device)
            self.stamps.view(64, 4)[:, 0] = -1       # min-reduced entry (uint64 max)
        self.dp_ranks = list(dp_ranks)
        n = len(params)
        self._sizes = (ctypes.c_longlong * n)(*[p.numel() for p in params])
        self._p = (ctypes.c_void_p * n)(*[p.data_ptr() for p in params])
        self._m = (ctypes.c_void_p * n)(*[m.data_ptr() for m in momentum_buffers])
        self._flags = (ctypes.c_void_p * self.world)(*[self.buf.ptr(r, self.flags_off) for r in range(self.world)])
        self._stages = [(ctypes.c_void_p * self.world)(*[self.buf.ptr(r, par * self.stage_bytes)
                                                        for r in range(self.world)]) for par in (0, 1)]
        self._ll = (ctypes.c_void_p * self.world)(*[self.buf.ptr(r, self.ll_off) for r in range(self.world)])

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0, zero_grad: bool = True) -> None:
        grads = self.grads_fn()
        n = len(self.params)
        g = (ctypes.c_void_p * n)(*[t.data_ptr() for t in grads])
        self.epoch += 1
        if self.protocol == "ll":
            rc = self.buf._lib.sdml_dp_ll_allreduce_sgd(
                n, self._p, g, self._m, self._sizes, self.index, self.world, self._ll,
                ctypes.c_void_p(self.epoch_dev.data_ptr()), ctypes.c_void_p(self.epoch_dev.data_ptr() + 4),
                float(self.lr), float(self.momentum), float(self.weight_decay), float(grad_scale),
                self.timeout_ns, ctypes.c_void_p(self.err_host.data_ptr()),
                ctypes.c_void_p(self.stamps.data_ptr()) if self.stamps is not None else None,
                torch.cuda.current_stream().cuda_stream)
            _lib.check(rc, "sdml_dp_ll_allreduce_sgd")
            return
        rc = self.buf._lib.sdml_dp_allreduce_sgd_dev_epoch(
            n, self._p, g, self._m, self._sizes, self.index, self.world, self._stages[1], self._stages[0],
            self._flags, ctypes.c_void_p(self.buf.ptr(self.index, self.counter_off)),
            ctypes.c_void_p(self.epoch_dev.data_ptr()), float(self.lr), float(self.momentum),
            float(self.weight_decay), float(grad_scale), torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_dp_allreduce_sgd_dev_epoch")
        self.epoch_dev[:1].add_(1)

    def check(self) -> None:
        """Raise if a poll of the low-latency kernel gave up on a replica (pinned host word: free)."""
        if int(self.err_host[0]) != 0:
            peer = self.dp_ranks[int(self.err_host[1])] if 0 <= int(self.err_host[1]) < self.world else "?"
            raise RuntimeError(f"data-parallel gradient exchange: gave up after {self.timeout_ns / 1e9:.0f}s waiting for "
                               f"replica rank {peer} at step {int(self.err_host[2]) & 0xFFFFFFFF}: it is dead or stalled")
