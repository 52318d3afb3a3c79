"""Python face of the C++ schedule generator (``csrc/schedule.cpp``).

The reference runs one batch at a time in lock-step across its two ranks
(/root/reference/simple_distributed.py:108-113); here a batch is split into
micro-batches that flow through GPipe / 1F1B / zero-bubble schedules.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from enum import IntEnum
from typing import List, Sequence

from ..ops import _lib


class OpKind(IntEnum):
    F = 0    # forward
    B = 1    # backward w.r.t. the stage input (dgrad)
    W = 2    # backward w.r.t. the weights (wgrad)
    BW = 3   # fused B + W


@dataclass(frozen=True)
class Op:
    kind: OpKind
    mb: int

    def __repr__(self) -> str:
        return f"{self.kind.name}{self.mb}"


@dataclass(frozen=True)
class Costs:
    """Relative op costs for the zero-bubble heuristic and the simulator."""
    f: float = 1.0
    b: float = 1.0
    w: float = 1.0
    c: float = 0.0


def _lib_sched() -> ctypes.CDLL:
    lib = _lib.sched_lib()
    if not getattr(lib, "_sdml_declared", False):
        D, I = ctypes.c_double, ctypes.c_int
        lib.sdml_sched_generate.restype = I
        lib.sdml_sched_generate.argtypes = [ctypes.c_char_p, I, I, I, D, D, D, D,
                                            ctypes.POINTER(I), I]
        lib.sdml_sched_simulate.restype = I
        lib.sdml_sched_simulate.argtypes = [ctypes.c_char_p, I, I, D, D, D, D,
                                            ctypes.POINTER(D), ctypes.POINTER(D), ctypes.POINTER(I)]
        lib.sdml_sched_validate.restype = I
        lib.sdml_sched_validate.argtypes = [ctypes.POINTER(I), I, I]
        lib.sdml_sched_generate_interleaved.restype = I
        lib.sdml_sched_generate_interleaved.argtypes = [I, I, I, I, I, D, D, D, D, ctypes.POINTER(I), I]
        lib.sdml_sched_simulate_interleaved.restype = I
        lib.sdml_sched_simulate_interleaved.argtypes = [I, I, I, I, D, D, D, D, ctypes.POINTER(D),
                                                        ctypes.POINTER(D), ctypes.POINTER(I)]
        lib._sdml_declared = True
    return lib


def generate(kind: str, p: int, m: int, rank: int, costs: Costs = Costs()) -> List[Op]:
    """Op list for ``rank`` of a ``p``-stage pipeline with ``m`` micro-batches."""
    lib = _lib_sched()
    max_ops = 3 * m + 8
    buf = (ctypes.c_int * (2 * max_ops))()
    n = lib.sdml_sched_generate(kind.encode(), p, m, rank, costs.f, costs.b, costs.w, costs.c,
                                buf, max_ops)
    if n < 0:
        raise ValueError(f"schedule generation failed (kind={kind!r}, p={p}, m={m}, rank={rank}): code {n}")
    return [Op(OpKind(buf[2 * i]), buf[2 * i + 1]) for i in range(n)]


def generate_all(kind: str, p: int, m: int, costs: Costs = Costs()) -> List[List[Op]]:
    return [generate(kind, p, m, r, costs) for r in range(p)]


@dataclass(frozen=True)
class SimResult:
    makespan: float
    busy: List[float]
    peak_inflight: List[int]

    @property
    def bubble_fraction(self) -> float:
        """1 - (busy time of the busiest-possible stage) / makespan, worst rank."""
        return max(1.0 - b / self.makespan for b in self.busy)

    @property
    def mean_bubble_fraction(self) -> float:
        return 1.0 - sum(self.busy) / (len(self.busy) * self.makespan)


def simulate(kind: str, p: int, m: int, costs: Costs = Costs()) -> SimResult:
    lib = _lib_sched()
    makespan = ctypes.c_double()
    busy = (ctypes.c_double * p)()
    peak = (ctypes.c_int * p)()
    rc = lib.sdml_sched_simulate(kind.encode(), p, m, costs.f, costs.b, costs.w, costs.c,
                                 ctypes.byref(makespan), busy, peak)
    if rc != 0:
        raise ValueError(f"schedule simulation failed (kind={kind!r}, p={p}, m={m}): code {rc}")
    return SimResult(makespan.value, list(busy), list(peak))


def validate(ops: Sequence[Op], m: int) -> None:
    lib = _lib_sched()
    flat = (ctypes.c_int * (2 * len(ops)))()
    for i, op in enumerate(ops):
        flat[2 * i] = int(op.kind)
        flat[2 * i + 1] = op.mb
    rc = lib.sdml_sched_validate(flat, len(ops), m)
    if rc != 0:
        raise ValueError(f"invalid schedule (code {rc}): {list(ops)}")


def splits_backward(kind: str) -> bool:
    return kind.startswith("zb")


# ---- interleaved (virtual-stage) zero-bubble schedules --------------------------------------------
@dataclass(frozen=True)
class ChunkOp:
    """One op of an interleaved schedule: ``chunk`` c of rank r is virtual stage c * p + r."""
    kind: OpKind
    mb: int
    chunk: int

    def __repr__(self) -> str:
        return f"{self.kind.name}{self.mb}.{self.chunk}"


def generate_interleaved(p: int, v: int, m: int, rank: int, costs: Costs = Costs(),
                         limit_chunks: int = 0) -> List[ChunkOp]:
    """Op list of ``rank`` when every rank holds ``v`` model chunks (costs are per chunk op)."""
    lib = _lib_sched()
    max_ops = 3 * m * v + 8
    buf = (ctypes.c_int * (3 * max_ops))()
    n = lib.sdml_sched_generate_interleaved(p, v, m, rank, limit_chunks, costs.f, costs.b, costs.w, costs.c,
                                            buf, max_ops)
    if n < 0:
        raise ValueError(f"interleaved schedule generation failed (p={p}, v={v}, m={m}, rank={rank}): code {n}")
    return [ChunkOp(OpKind(buf[3 * i]), buf[3 * i + 1], buf[3 * i + 2]) for i in range(n)]


def simulate_interleaved(p: int, v: int, m: int, costs: Costs = Costs(), limit_chunks: int = 0) -> SimResult:
    lib = _lib_sched()
    makespan = ctypes.c_double()
    busy = (ctypes.c_double * p)()
    peak = (ctypes.c_int * p)()
    rc = lib.sdml_sched_simulate_interleaved(p, v, m, limit_chunks, costs.f, costs.b, costs.w, costs.c,
                                             ctypes.byref(makespan), busy, peak)
    if rc != 0:
        raise ValueError(f"interleaved schedule simulation failed (p={p}, v={v}, m={m}): code {rc}")
    return SimResult(makespan.value, list(busy), list(peak))


def validate_interleaved(ops: Sequence[ChunkOp], m: int, v: int) -> None:
    """Every (chunk, micro-batch) has exactly one F, one B and one W in F < B < W order, and on every
    chunk the F's and the B's ascend (so sends and receives of a boundary pair up by order)."""
    pos = {}
    last = {}
    for i, op in enumerate(ops):
        if not (0 <= op.mb < m and 0 <= op.chunk < v) or op.kind not in (OpKind.F, OpKind.B, OpKind.W):
            raise ValueError(f"bad op {op}")
        key = (op.kind, op.chunk, op.mb)
        if key in pos:
            raise ValueError(f"duplicate {op}")
        pos[key] = i
        if op.kind != OpKind.W:
            if op.mb != last.get((op.kind, op.chunk), -1) + 1:
                raise ValueError(f"{op} out of order on its chunk")
            last[(op.kind, op.chunk)] = op.mb
    for c in range(v):
        for i in range(m):
            try:
                f, b, w = pos[(OpKind.F, c, i)], pos[(OpKind.B, c, i)], pos[(OpKind.W, c, i)]
            except KeyError as e:
                raise ValueError(f"missing op for chunk {c}, micro-batch {i}: {e}") from e
            if not f < b < w:
                raise ValueError(f"chunk {c}, micro-batch {i}: F/B/W order violated")
