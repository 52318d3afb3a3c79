"""Python face of ``csrc/launch_seq.cu``: a step's host-side CUDA calls (stream waits, event
records, async copies, a graph launch) as ONE native call.

Built once per (batch size, input-set parity); per step only the host source pointers and the loss
slot are patched.  No counterpart in the reference, whose loop blocks in an RPC per statement
(/root/reference/simple_distributed.py:108-113).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..ops import _lib

WAIT_EVENT, RECORD_EVENT, MEMCPY, GRAPH = 0, 1, 2, 3


class _Op(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int), ("flags", ctypes.c_int), ("a", ctypes.c_void_p),
                ("b", ctypes.c_void_p), ("c", ctypes.c_void_p), ("bytes", ctypes.c_ulonglong)]


def _declare(lib) -> None:
    if getattr(lib, "_sdml_launch_declared", False):
        return
    I, P = ctypes.c_int, ctypes.c_void_p
    lib.sdml_event_create.restype = I
    lib.sdml_event_create.argtypes = [ctypes.POINTER(P)]
    for name in ("sdml_event_destroy", "sdml_event_query", "sdml_event_synchronize"):
        fn = getattr(lib, name)
        fn.restype = I
        fn.argtypes = [P]
    lib.sdml_launch_seq.restype = I
    lib.sdml_launch_seq.argtypes = [ctypes.POINTER(_Op), I]
    lib._sdml_launch_declared = True


class NativeEvent:
    """A timing-disabled CUDA event owned by the native library."""

    def __init__(self):
        self.lib = _lib.cuda_lib()
        _declare(self.lib)
        h = ctypes.c_void_p()
        _lib.check(self.lib.sdml_event_create(ctypes.byref(h)), "sdml_event_create")
        self.handle = h.value

    def synchronize(self) -> None:
        _lib.check(self.lib.sdml_event_synchronize(self.handle), "sdml_event_synchronize")

    def done(self) -> bool:
        rc = self.lib.sdml_event_query(self.handle)
        if rc < 0:
            _lib.check(rc, "sdml_event_query")
        return rc == 0

    def __del__(self):
        try:
            if self.handle:
                self.lib.sdml_event_destroy(self.handle)
        except Exception:
            pass


class LaunchSeq:
    """An ordered list of primitive ops; ``run()`` issues them with one call into C++."""

    def __init__(self):
        self.lib = _lib.cuda_lib()
        _declare(self.lib)
        self._ops: List[_Op] = []
        self._arr = None
        self._keep: list = []          # objects whose handles are baked into the sequence

    def _add(self, kind: int, a, b, c=None, nbytes: int = 0, keep=()) -> int:
        assert self._arr is None, "sequence already frozen"
        self._ops.append(_Op(kind, 0, a or None, b or None, c or None, nbytes))
        self._keep.extend(keep)
        return len(self._ops) - 1

    def wait_event(self, stream: int, event: NativeEvent) -> int:
        return self._add(WAIT_EVENT, stream, event.handle, keep=(event,))

    def record_event(self, stream: int, event: NativeEvent) -> int:
        return self._add(RECORD_EVENT, stream, event.handle, keep=(event,))

    def memcpy(self, stream: int, dst: Optional[torch.Tensor], src: Optional[torch.Tensor], nbytes: int) -> int:
        return self._add(MEMCPY, stream, dst.data_ptr() if dst is not None else 0,
                         src.data_ptr() if src is not None else 0, nbytes, keep=(dst, src))

    def graph(self, stream: int, graph: torch.cuda.CUDAGraph) -> int:
        return self._add(GRAPH, stream, graph.raw_cuda_graph_exec(), keep=(graph,))

    def freeze(self) -> "LaunchSeq":
        self._arr = (_Op * len(self._ops))(*self._ops)
        return self

    # ---- per-call patches (indices returned by the builders above) ----
    def set_src(self, index: int, src: torch.Tensor) -> None:
        self._arr[index].c = src.data_ptr()

    def set_dst(self, index: int, dst: torch.Tensor) -> None:
        self._arr[index].b = dst.data_ptr()

    def set_event(self, index: int, event: NativeEvent) -> None:
        self._arr[index].b = event.handle

    def set_skip(self, index: int, skip: bool) -> None:
        self._arr[index].flags = 1 if skip else 0

    def run(self) -> None:
        rc = self.lib.sdml_launch_seq(self._arr, len(self._ops))
        if rc != 0:
            _lib.check(rc, "sdml_launch_seq")
