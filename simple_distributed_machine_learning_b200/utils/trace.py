"""Per-rank timeline of the pipeline executor: one span per scheduled op.

The reference has no profiler, timer or NVTX range anywhere
(/root/reference/simple_distributed.py: no ``time.`` / ``profil`` at all).
Here every F / B / W op, the optimizer step and the step as a whole can be bracketed:

* on a GPU by a pair of CUDA events recorded on the compute stream (the start event goes in
  AFTER the stream has been told to wait for the op's incoming p2p transfer, so a span is
  the op's own kernels, not the time it sat waiting for its neighbour) plus an NVTX range
  around the host-side launch code, so an external profiler shows the same names;
* on a CPU by ``time.perf_counter``.

From the spans: measured busy time, measured pipeline bubble (1 - busy / step period),
per-op-kind means, and a Chrome-trace (``chrome://tracing`` / Perfetto) JSON file per rank.
Nothing is recorded unless a tracer is attached to the engine; event pairs are resolved
lazily (one device synchronise in ``resolve()``), never in the hot loop.
"""
from __future__ import annotations

import json
import os
import time
from collections import defaultdict
from contextlib import contextmanager
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch


@dataclass
class Span:
    step: int
    kind: str          # "F" | "B" | "W" | "BW" | "opt" | "step"
    mb: int            # micro-batch, -1 when not applicable
    start_ms: float
    end_ms: float

    @property
    def dur_ms(self) -> float:
        return self.end_ms - self.start_ms


class StepTracer:
    """Collects spans for at most ``max_steps`` steps (after skipping ``skip`` warm-up steps)."""

    def __init__(self, device: torch.device, rank: int = 0, stage: int = 0, skip: int = 0,
                 max_steps: int = 64, nvtx: bool = True):
        self.device = device
        self.rank, self.stage = rank, stage
        self.skip, self.max_steps = skip, max_steps
        self.is_cuda = device.type == "cuda"
        self.nvtx = bool(nvtx and self.is_cuda)
        self._raw: List[tuple] = []       # (step, kind, mb, start, end) with events or floats
        self._spans: List[Span] = []
        self._t0 = None
        self._step = -1
        self._seen = 0
        self._active = False

    # ------------------------------------------------------------------ recording
    def _now(self):
        if self.is_cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
            return ev
        return time.perf_counter()

    def begin_step(self) -> None:
        self._seen += 1
        self._active = self.skip < self._seen <= self.skip + self.max_steps
        if not self._active:
            return
        self._step += 1
        self._step_start = self._now()
        if self._t0 is None:
            self._t0 = self._step_start

    def end_step(self) -> None:
        if self._active:
            self._raw.append((self._step, "step", -1, self._step_start, self._now()))
            self._active = False

    @contextmanager
    def span(self, kind: str, mb: int = -1):
        if not self._active:
            yield
            return
        if self.nvtx:
            torch.cuda.nvtx.range_push(f"{kind}{mb if mb >= 0 else ''} s{self.stage}")
        start = self._now()
        try:
            yield
        finally:
            self._raw.append((self._step, kind, mb, start, self._now()))
            if self.nvtx:
                torch.cuda.nvtx.range_pop()

    # ------------------------------------------------------------------ read-out
    def resolve(self) -> List[Span]:
        """Turn recorded event pairs into milliseconds since the first traced step began."""
        if self._raw:
            if self.is_cuda:
                torch.cuda.synchronize(self.device)
                conv = lambda e: self._t0.elapsed_time(e)
            else:
                conv = lambda t: (t - self._t0) * 1e3
            self._spans.extend(Span(s, k, mb, conv(a), conv(b)) for s, k, mb, a, b in self._raw)
            self._raw.clear()
        return self._spans

    def summary(self) -> Dict[str, object]:
        """Measured busy / bubble figures of this rank.

        ``step_ms`` is the steady-state period (start of one traced step to the start of the
        next; the span of the single step when only one was traced), ``busy_ms`` the mean
        summed duration of the ops inside a step, ``bubble`` = 1 - busy / period."""
        spans = self.resolve()
        steps = sorted((s for s in spans if s.kind == "step"), key=lambda s: s.step)
        if not steps:
            return {"rank": self.rank, "stage": self.stage, "steps": 0}
        if len(steps) > 1:
            period = (steps[-1].start_ms - steps[0].start_ms) / (len(steps) - 1)
            counted = {s.step for s in steps[:-1]}
        else:
            period = steps[0].dur_ms
            counted = {steps[0].step}
        busy = defaultdict(float)
        per_kind = defaultdict(list)
        for s in spans:
            if s.kind == "step":
                continue
            per_kind[s.kind].append(s.dur_ms)
            if s.step in counted:
                busy[s.step] += s.dur_ms
        busy_ms = sum(busy.values()) / max(1, len(counted))
        return {
            "rank": self.rank, "stage": self.stage, "steps": len(steps),
            "step_ms": period, "busy_ms": busy_ms,
            "bubble": max(0.0, 1.0 - busy_ms / period) if period > 0 else 0.0,
            "mean_ms": {k: sum(v) / len(v) for k, v in sorted(per_kind.items())},
            "clock": "cuda_events" if self.is_cuda else "perf_counter",
        }

    def chrome_trace(self) -> List[dict]:
        """Complete-event ("ph": "X") records; pid = pipeline stage, tid = rank."""
        out = [{"name": "process_name", "ph": "M", "pid": self.stage, "tid": self.rank,
                "args": {"name": f"stage {self.stage}"}},
               {"name": "thread_name", "ph": "M", "pid": self.stage, "tid": self.rank,
                "args": {"name": f"rank {self.rank}"}}]
        for s in self.resolve():
            if s.kind == "step":
                continue
            out.append({"name": f"{s.kind}{s.mb}" if s.mb >= 0 else s.kind, "cat": s.kind, "ph": "X",
                        "pid": self.stage, "tid": self.rank, "ts": s.start_ms * 1e3,
                        "dur": max(s.dur_ms, 0.0) * 1e3, "args": {"step": s.step, "mb": s.mb}})
        return out

    def dump(self, trace_dir: str) -> str:
        os.makedirs(trace_dir, exist_ok=True)
        path = os.path.join(trace_dir, f"trace_rank{self.rank}.json")
        with open(path, "w") as f:
            json.dump({"traceEvents": self.chrome_trace(), "displayTimeUnit": "ms",
                       "summary": self.summary()}, f)
        return path


def merge_traces(trace_dir: str, out_name: str = "trace_all.json") -> Optional[str]:
    """Concatenate the per-rank files of ``trace_dir`` into one viewable file (rank clocks are
    not aligned with each other: every rank's time axis starts at its own first traced step)."""
    events, summaries = [], []
    for name in sorted(os.listdir(trace_dir)):
        if name.startswith("trace_rank") and name.endswith(".json"):
            with open(os.path.join(trace_dir, name)) as f:
                d = json.load(f)
            events.extend(d["traceEvents"])
            summaries.append(d.get("summary"))
    if not events:
        return None
    path = os.path.join(trace_dir, out_name)
    with open(path, "w") as f:
        json.dump({"traceEvents": events, "displayTimeUnit": "ms", "summaries": summaries}, f)
    return path
