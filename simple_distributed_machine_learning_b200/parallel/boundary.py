"""Stage-boundary transport over symmetric memory (``csrc/boundary.cu``): push + flag, no NCCL.

The reference hands an activation to the next rank with a blocking RPC and gets the gradient
back the same way (/root/reference/simple_distributed.py:47-49, :71, :80, :112).  Here every
pipeline rank owns one peer-mapped region (``csrc/symm.cu``) that holds its RECEIVE slots:

    [header: flags + acknowledgement words][activations in][gradients in][eval activations in]

and maps the regions of its two neighbours.  A sender writes the neighbour's slot directly over
NVLink -- the producing GEMM's TMA-store epilogue when the stage supports it (``MLPStage``), or
one small copy kernel -- then publishes a 4-byte flag; the receiver's stream waits for the flag
with a one-thread kernel in front of the consumer.  Everything is stream-ordered work on the
compute stream, so a whole training step of a pipeline rank, transfers included, is capturable
into one CUDA graph (round counters live on the device).

Slot reuse is guarded by one acknowledgement per direction and round; every wait is bounded by
``timeout_s`` and reports through :meth:`SymmBoundary.check` (the reference disables every
timeout, ref:36 / ref:167).
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib
from .symm import SymmetricBuffer

_HDR = 4096
_MAX_SLOTS = 256
_OFF_ACT_FLAG = 0            # uint32[_MAX_SLOTS]  written by the previous stage
_OFF_GRAD_FLAG = 1024        # uint32[_MAX_SLOTS]  written by the next stage
_OFF_ACT_ACK = 2048          # written by the NEXT stage: it has consumed my activations of round r
_OFF_GRAD_ACK = 2052         # written by the PREVIOUS stage: it has consumed my gradients of round r
_OFF_EVAL_FLAG = 2056        # written by the previous stage (evaluation batches)
_OFF_EVAL_ACK = 2060         # written by the next stage

# error codes a timed-out wait leaves behind
_CODES = {1: "activation", 2: "gradient", 3: "acknowledgement of the activations", 4: "acknowledgement of the gradients",
          5: "evaluation activation", 6: "acknowledgement of the evaluation activations"}


class BoundaryTimeout(RuntimeError):
    pass


# ---- raw device memory as a torch tensor on THIS device (DLPack capsule built by hand) ----------
# ``__cuda_array_interface__`` lets torch look the pointer up (cudaPointerGetAttributes), and a
# peer-mapped pointer reports the PEER's device; a DLPack capsule states the device itself.
class _DLDevice(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int32), ("device_id", ctypes.c_int32)]


class _DLDataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class _DLTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("device", _DLDevice), ("ndim", ctypes.c_int32),
                ("dtype", _DLDataType), ("shape", ctypes.POINTER(ctypes.c_int64)),
                ("strides", ctypes.POINTER(ctypes.c_int64)), ("byte_offset", ctypes.c_uint64)]


class _DLManagedTensor(ctypes.Structure):
    pass


_DELETER = ctypes.CFUNCTYPE(None, ctypes.POINTER(_DLManagedTensor))
_DLManagedTensor._fields_ = [("dl_tensor", _DLTensor), ("manager_ctx", ctypes.c_void_p), ("deleter", _DELETER)]
_noop_deleter = _DELETER(lambda _p: None)
_dl_keep: List[object] = []      # the structs outlive the tensors (a few hundred bytes per view, never freed)


def device_bytes(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    """uint8 tensor of ``nbytes`` at device address ``ptr``, declared to live on ``device``."""
    shape = (ctypes.c_int64 * 1)(nbytes)
    mt = _DLManagedTensor()
    mt.dl_tensor.data = ctypes.c_void_p(ptr)
    mt.dl_tensor.device = _DLDevice(2, device.index if device.index is not None else torch.cuda.current_device())
    mt.dl_tensor.ndim = 1
    mt.dl_tensor.dtype = _DLDataType(1, 8, 1)      # kDLUInt, 8 bits
    mt.dl_tensor.shape = shape
    mt.dl_tensor.strides = None
    mt.dl_tensor.byte_offset = 0
    mt.manager_ctx = None
    mt.deleter = _noop_deleter
    _dl_keep.append((mt, shape))
    new = ctypes.pythonapi.PyCapsule_New
    new.restype = ctypes.py_object
    new.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
    cap = new(ctypes.addressof(mt), b"dltensor", None)
    return torch.from_dlpack(cap)


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


def _numel(shape: Sequence[int]) -> int:
    n = 1
    for s in shape:
        n *= int(s)
    return n


class SymmBoundary:
    """Receive slots, flags and peer mappings of ONE pipeline rank.  Built collectively: every rank
    of the job constructs its instance at the same point (the IPC handles and region layouts
    travel over the control group)."""

    def __init__(self, topo, rank: int, ctrl_group, device: torch.device, stage, max_rows: int,
                 timeout_s: float = 300.0):
        self.lib = _lib.cuda_lib()
        self.device, self.rank = device, rank
        self.prev, self.next = topo.prev_stage_rank(rank), topo.next_stage_rank(rank)
        self.timeout_ns = int(max(0.0, timeout_s) * 1e9)
        self.dtype = stage.act_dtype
        self.esz = torch.empty((), dtype=self.dtype).element_size()
        in_row = _numel(stage.in_shape(1)) * self.esz if self.prev is not None else 0
        out_row = _numel(stage.out_shape(1)) * self.esz if self.next is not None else 0
        self.in_row_bytes, self.out_row_bytes = in_row, out_row
        self.max_rows = max_rows
        act_b, grad_b = _align(max_rows * in_row), _align(max_rows * out_row)
        self.layout = {"act": _HDR, "grad": _HDR + act_b, "eval": _HDR + act_b + grad_b,
                       "in_row": in_row, "out_row": out_row}
        total = _HDR + 2 * act_b + grad_b
        ranks = sorted({r for r in (self.prev, rank, self.next) if r is not None})
        self.buf = SymmetricBuffer(ranks, rank, total, ctrl_group, device)
        world = dist.get_world_size(group=ctrl_group)
        gathered: List[Optional[tuple]] = [None] * world
        dist.all_gather_object(gathered, (rank, self.layout), group=ctrl_group)
        layouts = {r: l for (r, l) in gathered}
        self._base = {r: self.buf.ptrs[i] for i, r in enumerate(ranks)}
        self._peer_layout = {r: layouts[r] for r in ranks}
        for peer, mine, theirs in ((self.next, "out_row", "in_row"), (self.prev, "in_row", "out_row")):
            if peer is not None and self.layout[mine] != layouts[peer][theirs]:
                raise ValueError(f"rank {rank}: boundary row size {self.layout[mine]} B does not match "
                                 f"rank {peer}'s {layouts[peer][theirs]} B")
        # round counters (device): [0] training steps completed, [1] evaluation batches completed
        self.cnt = torch.zeros(2, dtype=torch.int32, device=device)
        self.err_dev = torch.zeros(4, dtype=torch.int32, device=device)
        self.err_host = torch.zeros(4, dtype=torch.int32).pin_memory()
        self._offs: List[int] = [0]
        self._acked = {"act": False, "grad": False}
        self._views: Dict[tuple, torch.Tensor] = {}
        self.direct_writes = 0
        self.pushes = 0

    # ---- addresses ----
    def _hdr(self, r: int, off: int) -> int:
        return self._base[r] + off

    def _region(self, r: int, kind: str) -> int:
        return self._base[r] + self._peer_layout[r][kind]

    def _view(self, r: int, kind: str, row0: int, rows: int, row_shape: Tuple[int, ...], row_bytes: int) -> torch.Tensor:
        key = (r, kind, row0, rows, row_shape)
        t = self._views.get(key)
        if t is None:
            if row0 + rows > self.max_rows:
                raise ValueError(f"boundary region holds {self.max_rows} rows, asked for rows [{row0}, {row0 + rows})")
            raw = device_bytes(self._region(r, kind) + row0 * row_bytes, rows * row_bytes, self.device)
            t = self._views[key] = raw.view(self.dtype).view(rows, *row_shape)
        return t

    # ---- round set-up ----
    def begin_round(self, sizes: Sequence[int]) -> None:
        offs = [0]
        for s in sizes:
            offs.append(offs[-1] + int(s))
        if len(sizes) > _MAX_SLOTS:
            raise ValueError(f"at most {_MAX_SLOTS} micro-batches per step on the symmetric boundary")
        self._offs = offs
        self._acked = {"act": False, "grad": False}

    # ---- slots ----
    def in_slot(self, kind: str, mb: int, row_shape: Tuple[int, ...]) -> torch.Tensor:
        """LOCAL receive slot of micro-batch ``mb`` ("act" from the previous stage, "grad" from the next)."""
        rb = self.in_row_bytes if kind == "act" else self.out_row_bytes
        return self._view(self.rank, kind, self._offs[mb], self._offs[mb + 1] - self._offs[mb], tuple(row_shape), rb)

    def in_region(self, kind: str, rows: int, row_shape: Tuple[int, ...]) -> torch.Tensor:
        """All local receive rows of a step as one row-stacked tensor (``MLPStage`` layout)."""
        rb = self.in_row_bytes if kind in ("act", "eval") else self.out_row_bytes
        return self._view(self.rank, kind, 0, rows, tuple(row_shape), rb)

    def out_slot(self, kind: str, mb: int, row_shape: Tuple[int, ...]) -> torch.Tensor:
        """The NEIGHBOUR's receive slot of micro-batch ``mb``, mapped here: a kernel that writes its
        output to this tensor has sent it ("act": next stage, "grad": previous stage).  Must be
        followed by :meth:`send` of the same tensor, which publishes the flag."""
        peer = self.next if kind == "act" else self.prev
        rb = self.out_row_bytes if kind == "act" else self.in_row_bytes
        self._ensure_ack(kind)
        return self._view(peer, kind, self._offs[mb], self._offs[mb + 1] - self._offs[mb], tuple(row_shape), rb)

    # ---- protocol steps (all enqueued on the current stream) ----
    def _s(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _cnt_ptr(self, which: int) -> int:
        return self.cnt.data_ptr() + 4 * which

    def _wait(self, flag_ptr: int, which: int, add: int, code: int) -> None:
        _lib.check(self.lib.sdml_bnd_wait(flag_ptr, self._cnt_ptr(which), add, self.timeout_ns,
                                          self.err_dev.data_ptr(), self.err_host.data_ptr(), code, self._s()),
                   "sdml_bnd_wait")

    def _signal(self, flag_ptr: Optional[int], which: int, add: int, bump: int) -> None:
        _lib.check(self.lib.sdml_bnd_signal(flag_ptr, self._cnt_ptr(which), add, bump, self._s()), "sdml_bnd_signal")

    def _ensure_ack(self, kind: str) -> None:
        """Before the first write of a round into the neighbour's slots: it must have finished reading
        them in the previous round."""
        if self._acked[kind]:
            return
        self._acked[kind] = True
        off, code = (_OFF_ACT_ACK, 3) if kind == "act" else (_OFF_GRAD_ACK, 4)
        self._wait(self._hdr(self.rank, off), 0, 0, code)

    def wait(self, kind: str, mb: int) -> None:
        """The current stream continues once micro-batch ``mb``'s slot has been filled this round."""
        off, code = (_OFF_ACT_FLAG, 1) if kind == "act" else (_OFF_GRAD_FLAG, 2)
        self._wait(self._hdr(self.rank, off + 4 * mb), 0, 1, code)

    def send(self, kind: str, mb: int, t: torch.Tensor) -> None:
        """Deliver ``t`` as micro-batch ``mb`` to the neighbour.  If ``t`` IS the neighbour's slot
        (:meth:`out_slot`), the data is already there and only the flag is published."""
        peer = self.next if kind == "act" else self.prev
        rb = self.out_row_bytes if kind == "act" else self.in_row_bytes
        flag = self._hdr(peer, (_OFF_ACT_FLAG if kind == "act" else _OFF_GRAD_FLAG) + 4 * mb)
        dst = self._region(peer, kind) + self._offs[mb] * rb
        if t.data_ptr() == dst:
            self.direct_writes += 1
            self._signal(flag, 0, 1, 0)
            return
        nbytes = t.numel() * t.element_size()
        assert t.is_contiguous() and nbytes == (self._offs[mb + 1] - self._offs[mb]) * rb, "boundary tensor shape"
        ack = None
        if not self._acked[kind]:
            self._acked[kind] = True
            ack = self._hdr(self.rank, _OFF_ACT_ACK if kind == "act" else _OFF_GRAD_ACK)
        self.pushes += 1
        _lib.check(self.lib.sdml_bnd_push(dst, t.data_ptr(), nbytes, flag, self._cnt_ptr(0), 1, ack, self.timeout_ns,
                                          self.err_dev.data_ptr(), self.err_host.data_ptr(),
                                          3 if kind == "act" else 4, self._s()), "sdml_bnd_push")

    def end_round(self) -> None:
        """After the last reader of this round's slots has been enqueued: acknowledge both neighbours
        and advance the round counter."""
        if self.prev is not None:
            self._signal(self._hdr(self.prev, _OFF_ACT_ACK), 0, 1, 0 if self.next is not None else 1)
        if self.next is not None:
            self._signal(self._hdr(self.next, _OFF_GRAD_ACK), 0, 1, 1)

    # ---- evaluation batches: one slot, acknowledged per batch ----
    def eval_recv(self, rows: int, row_shape: Tuple[int, ...]) -> torch.Tensor:
        self._wait(self._hdr(self.rank, _OFF_EVAL_FLAG), 1, 1, 5)
        return self.in_region("eval", rows, row_shape)

    def eval_send(self, y: torch.Tensor) -> None:
        nbytes = y.numel() * y.element_size()
        assert y.is_contiguous() and nbytes <= self.max_rows * self.out_row_bytes
        pad = (-nbytes) % 16
        assert pad == 0, "evaluation activations must be a multiple of 16 bytes"
        _lib.check(self.lib.sdml_bnd_push(self._region(self.next, "eval"), y.data_ptr(), nbytes,
                                          self._hdr(self.next, _OFF_EVAL_FLAG), self._cnt_ptr(1), 1,
                                          self._hdr(self.rank, _OFF_EVAL_ACK), self.timeout_ns,
                                          self.err_dev.data_ptr(), self.err_host.data_ptr(), 6, self._s()),
                   "sdml_bnd_push")

    def eval_end(self) -> None:
        self._signal(self._hdr(self.prev, _OFF_EVAL_ACK) if self.prev is not None else None, 1, 1, 1)

    def debug_set_round(self, value: int) -> None:
        """Tests only: restart the protocol at round ``value`` (counters, and every local flag /
        acknowledgement word to the same "completed rounds" figure).  Every rank calls it with the
        same value, followed by a barrier."""
        v = value & 0xFFFFFFFF
        v = v - (1 << 32) if v >= (1 << 31) else v
        self.cnt.fill_(v)
        hdr = self.buf.local_view(torch.int32, 0, _HDR // 4)
        hdr.fill_(v)
        torch.cuda.synchronize(self.device)

    # ---- failure detection ----
    def check(self) -> None:
        """Raise if a wait of this rank has timed out (host-side poll of pinned memory: free)."""
        code = int(self.err_host[0])
        if code == 0:
            return
        seen, want = int(self.err_host[1]) & 0xFFFFFFFF, int(self.err_host[2]) & 0xFFFFFFFF
        what = _CODES.get(code, f"code {code}")
        peer = {1: self.prev, 2: self.next, 3: self.next, 4: self.prev, 5: self.prev, 6: self.next}.get(code)
        raise BoundaryTimeout(
            f"rank {self.rank}: gave up after {self.timeout_ns / 1e9:.0f}s waiting for the {what} from rank {peer} "
            f"(flag {seen}, wanted {want}): the peer is dead or stalled")

    def close(self) -> None:
        self._views.clear()
        self.buf.close()
