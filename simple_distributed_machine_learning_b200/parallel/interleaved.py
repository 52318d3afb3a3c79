"""Interleaved (virtual-stage) pipeline execution: every rank holds ``v`` model chunks.

The reference has one batch in flight and exactly two stages (/root/reference/simple_distributed.py:47-50,
:108-113).  With one chunk per rank a single training step cannot idle less than
(p - 1) / (3m + p - 1) of the time; giving every rank ``v`` non-adjacent chunks --
chunk ``c`` of rank ``r`` is virtual stage ``q = c * p + r`` of ``p * v`` -- shrinks the ramp to chunk-sized
ops: ((p - 1) / v) / (3m + (p - 1) / v), e.g. 5.9 % instead of 11.1 % for p = 4 at 8 micro-batches.

This module is the executor for those schedules (``csrc/schedule.cpp: gen_zb_interleaved``):

* :class:`ChunkedStage` -- the ``v`` stages of a rank behind the few attributes the Trainer needs
  (loss, checkpoint parts keyed by VIRTUAL stage index, optimizer state);
* :class:`InterleavedEngine` -- walks the (kind, micro-batch, chunk) list; a micro-batch goes from virtual
  stage q to q + 1 by point-to-point send / receive on a communicator dedicated to that virtual boundary
  and direction (rank r -> r + 1, and p - 1 -> 0 between chunks), receives posted at their point of use.

Transport here is ``torch.distributed`` p2p (gloo on CPUs, NCCL on GPUs); the peer-mapped boundary and the
whole-step CUDA graph of the one-chunk executor (``parallel/pipeline.py``) are not wired to it yet.
"""
from __future__ import annotations

import datetime
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from . import schedule as sched
from .schedule import OpKind
from ..utils.data import split_microbatches


class _ChunkedOptimizer:
    """The chunks' optimizers as one object (state_dict / load_state_dict / step_count)."""

    def __init__(self, opts):
        self.opts = list(opts)

    @property
    def step_count(self) -> int:
        return self.opts[0].step_count

    @step_count.setter
    def step_count(self, v: int) -> None:
        for o in self.opts:
            o.step_count = v

    def state_dict(self) -> Dict:
        return {"kind": "chunked", "chunks": [o.state_dict() for o in self.opts]}

    def load_state_dict(self, sd: Dict) -> None:
        assert sd.get("kind") == "chunked" and len(sd["chunks"]) == len(self.opts)
        for o, s in zip(self.opts, sd["chunks"]):
            o.load_state_dict(s)


class ChunkedStage:
    """``chunks[c]`` is virtual stage ``c * p + pp_idx``; the rank needs inputs when it holds virtual
    stage 0 (rank 0) and owns the loss when it holds the last one (rank p - 1)."""

    supports_symm_boundary = False

    def __init__(self, chunks: List, pp_idx: int, p: int):
        self.chunks, self.pp_idx, self.p, self.v = list(chunks), pp_idx, p, len(chunks)
        self.device = chunks[0].device
        self.act_dtype = chunks[0].act_dtype
        self.is_first = chunks[0].is_first
        self.is_last = chunks[-1].is_last
        self.optimizer = _ChunkedOptimizer([c.optimizer for c in chunks])
        self.dp_group, self.dp_size = None, 1

    def q_of(self, chunk: int) -> int:
        return chunk * self.p + self.pp_idx

    # ---- what the Trainer reads ----
    @property
    def loss_sum(self) -> torch.Tensor:
        return self.chunks[-1].loss_sum

    @property
    def correct(self) -> torch.Tensor:
        return self.chunks[-1].correct

    def mean_loss(self) -> torch.Tensor:
        return self.chunks[-1].mean_loss()

    def set_training(self, training: bool) -> None:
        for c in self.chunks:
            c.set_training(training)

    def eval_begin(self) -> None:
        for c in self.chunks:
            c.eval_begin()

    def in_shape(self, n: int):
        return self.chunks[0].in_shape(n)

    def out_shape(self, n: int):
        return self.chunks[-1].out_shape(n)

    def flops_per_sample(self) -> int:
        return sum(c.flops_per_sample() for c in self.chunks)

    # ---- checkpoint: one file per VIRTUAL stage, same format as one-chunk pipelines of p * v stages ----
    def checkpoint_part_indices(self, pp_idx: int) -> List[int]:
        return [self.q_of(c) for c in range(self.v)]

    def checkpoint_parts(self, pp_idx: int) -> Dict[int, Dict[str, torch.Tensor]]:
        return {self.q_of(c): ch.state_dict() for c, ch in enumerate(self.chunks)}

    def load_checkpoint_parts(self, pp_idx: int, parts: Dict[int, Dict[str, torch.Tensor]]) -> None:
        for c, ch in enumerate(self.chunks):
            ch.load_state_dict(parts[self.q_of(c)])


class InterleavedEngine:
    """Runs one rank's interleaved op list.  Same step protocol as ``PipelineEngine`` (``train_step`` /
    ``eval_batch``), one process per rank, data-/tensor-parallel size 1."""

    def __init__(self, stage: ChunkedStage, topo, rank: int, microbatches: int, timeout_s: float = 300.0,
                 backend: Optional[str] = None):
        assert topo.dp == 1 and topo.tp == 1, "interleaved schedules: pipeline parallelism only"
        self.stage, self.topo, self.rank = stage, topo, rank
        self.p, self.v = topo.pp, stage.v
        self.P = self.p * self.v                      # virtual stages
        self.m = microbatches
        self.timeout_s = timeout_s
        self.is_cuda = stage.device.type == "cuda"
        self.steps_done = 0
        self.tracer = None
        # attributes the Trainer looks at on any engine
        self.bnd = None
        self.native = None
        self.executor = "python"
        self.p2p_mode = "torch"
        self._ops_cache: Dict[int, List[sched.ChunkOp]] = {}
        self._local: Dict[tuple, torch.Tensor] = {}   # hand-off between chunks of ONE rank (p == 1)
        self._sends: List[tuple] = []
        # one communicator per virtual boundary and direction; every rank creates all of them in the
        # same order (new_group is collective over the job)
        self.fwd: Dict[int, object] = {}
        self.bwd: Dict[int, object] = {}
        if self.p > 1:
            for b in range(self.P - 1):
                pair = sorted({b % self.p, (b + 1) % self.p})
                gf = dist.new_group(pair, backend=backend)
                gb = dist.new_group(pair, backend=backend)
                if rank in pair:
                    self.fwd[b], self.bwd[b] = gf, gb
            self._warmup()

    def _rank_of(self, q: int) -> int:
        return q % self.p

    def _warmup(self) -> None:
        """Open every connection once, in one global order (NCCL connects lazily inside the first
        send / receive and blocks the host until the peer shows up)."""
        token = torch.zeros(8, dtype=torch.float32, device=self.stage.device)
        for b in range(self.P - 1):
            up, down = self._rank_of(b), self._rank_of(b + 1)
            if self.rank == up:
                dist.send(token, dst=down, group=self.fwd[b])
                dist.recv(token, src=down, group=self.bwd[b])
            elif self.rank == down:
                dist.recv(token, src=up, group=self.fwd[b])
                dist.send(token, dst=up, group=self.bwd[b])
            if self.is_cuda:
                torch.cuda.synchronize(self.stage.device)

    def ops_for(self, n_mb: int) -> List[sched.ChunkOp]:
        ops = self._ops_cache.get(n_mb)
        if ops is None:
            ops = sched.generate_interleaved(self.p, self.v, n_mb, self.stage.pp_idx)
            sched.validate_interleaved(ops, n_mb, self.v)
            self._ops_cache[n_mb] = ops
        return ops

    # ---- transfers across virtual boundary b (between virtual stages b and b + 1) ----
    def _send(self, kind: str, b: int, mb: int, t: torch.Tensor) -> None:
        if self.p == 1:
            self._local[(kind, b, mb)] = t
            return
        t = t.contiguous()
        if kind == "act":
            w = dist.isend(t, dst=self._rank_of(b + 1), group=self.fwd[b])
        else:
            w = dist.isend(t, dst=self._rank_of(b), group=self.bwd[b])
        self._sends.append((w, t))

    def _recv(self, kind: str, b: int, mb: int, shape, buf: Optional[torch.Tensor]) -> torch.Tensor:
        if self.p == 1:
            return self._local.pop((kind, b, mb))
        if buf is None:
            buf = torch.empty(shape, dtype=self.stage.act_dtype, device=self.stage.device)
        if kind == "act":
            w = dist.irecv(buf, src=self._rank_of(b), group=self.fwd[b])
        else:
            w = dist.irecv(buf, src=self._rank_of(b + 1), group=self.bwd[b])
        self._wait(w, f"{kind} of micro-batch {mb} across virtual boundary {b}")
        return buf

    def _wait(self, w, what: str) -> None:
        from .pipeline import PipelineTimeout
        if self.is_cuda:
            w.wait()            # stream-side; a dead peer surfaces through the process group's watchdog
            return
        try:
            w.wait(datetime.timedelta(seconds=self.timeout_s))
        except RuntimeError as e:
            raise PipelineTimeout(f"rank {self.rank}: gave up after {self.timeout_s:.0f}s waiting for the {what}: {e}") from e

    def _drain(self) -> None:
        for w, _t in self._sends:
            self._wait(w, "completion of a send")
        self._sends.clear()

    # ---- one optimizer step ----
    def train_step(self, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor], batch_size: int):
        st = self.stage
        sizes = split_microbatches(batch_size, self.m)
        n_mb = len(sizes)
        offs = [0]
        for s in sizes:
            offs.append(offs[-1] + s)
        ops = self.ops_for(n_mb)
        for ch in st.chunks:
            ch.begin_step(batch_size, sizes)
        for op in ops:
            ch, q, mb = st.chunks[op.chunk], st.q_of(op.chunk), op.mb
            if op.kind == OpKind.F:
                if q == 0:
                    x = inputs[offs[mb]:offs[mb + 1]]
                else:
                    x = self._recv("act", q - 1, mb, ch.in_shape(sizes[mb]), ch.input_buffer(mb))
                t = targets[offs[mb]:offs[mb + 1]] if q == self.P - 1 else None
                y = ch.forward(mb, x, t)
                if q < self.P - 1:
                    self._send("act", q, mb, y)
            elif op.kind == OpKind.B:
                dy = None
                if q < self.P - 1:
                    dy = self._recv("grad", q, mb, ch.out_shape(sizes[mb]), ch.grad_buffer(mb))
                dx = ch.backward_input(mb, dy)
                if q > 0:
                    self._send("grad", q - 1, mb, dx)
            else:
                ch.backward_weight(mb)
        for ch in st.chunks:
            ch.end_step()
        self._drain()
        self.steps_done += 1
        return st.mean_loss() if st.is_last else None

    @torch.no_grad()
    def eval_batch(self, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor], batch_size: int) -> None:
        st = self.stage
        for c, ch in enumerate(st.chunks):
            q = st.q_of(c)
            x = inputs if q == 0 else self._recv("act", q - 1, 0, ch.in_shape(batch_size), None)
            y = ch.eval_forward(x, targets if q == self.P - 1 else None)
            if q < self.P - 1:
                self._send("act", q, 0, y)
        self._drain()
