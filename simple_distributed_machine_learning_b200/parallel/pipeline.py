"""Schedule executor: runs one rank's op list (F / B / W) for every training step.

Replaces the reference's hot loop (/root/reference/simple_distributed.py:108-113):
``forward`` -> blocking ``rpc_sync`` (ref:49) -> ``dist_autograd.backward`` (ref:112) ->
``dist_optim.step`` (ref:113), one batch in flight.  Here:

* the batch is split into micro-batches that follow a GPipe / 1F1B / zero-bubble op list
  produced by the C++ generator (``csrc/schedule.cpp``);
* on GPUs stage-boundary tensors are PUSHED into the neighbour's peer-mapped receive slot
  (``parallel/boundary.py`` / ``csrc/boundary.cu``: the producing GEMM's epilogue or one small
  copy kernel writes over NVLink, a flag follows, the consumer's stream waits for the flag) --
  no NCCL call, and the whole step of a rank is one capturable stream of kernels;
* otherwise (CPUs, ``--p2p torch|native``) they move with point-to-point sends/receives (NCCL over
  NVLink on GPUs, gloo on CPUs) on communicators dedicated to one boundary and one direction.
  A receive is enqueued (on its communicator's own stream) as soon as the previous op of
  this rank has been *launched* -- the host runs ahead of the device, so the transfer
  overlaps the previous op's kernels -- and sends never block the compute stream.
  Posting receives any earlier (``lookahead`` > 0) is unsafe on a GPU: a receive kernel
  that spins for a message whose sender is still waiting for work this rank has not
  launched yet can sit in front of that very work in a shared hardware queue (measured:
  lookahead 2 dead-locks 2 x B200, lookahead 0 does not);
* there is no barrier between steps: a rank starts the next step as soon as its own
  optimizer update is enqueued, so the pipeline ramp of step k+1 overlaps the drain of
  step k.

Failure handling: the reference disables every timeout (``rpc_timeout=0``, ref:167) and
hangs forever if the peer dies.  Here every wait is bounded by ``timeout_s`` and raises
``PipelineTimeout`` naming the peer, the direction and the micro-batch.
"""
from __future__ import annotations

import contextlib
import datetime
import os
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from . import schedule as sched
from .schedule import Op, OpKind
from .stage import Stage
from .topology import ProcessGroups, Topology
from ..utils.data import split_microbatches


class PipelineTimeout(RuntimeError):
    pass


@contextlib.contextmanager
def _no_span(kind: str, mb: int = -1):
    yield


@dataclass
class _Pending:
    work: object
    buf: torch.Tensor
    what: str


class PipelineEngine:
    def __init__(self, stage: Stage, topo: Topology, rank: int, groups: Optional[ProcessGroups],
                 microbatches: int = 1, schedule: str = "1f1b", timeout_s: float = 300.0,
                 lookahead: int = 0, costs: sched.Costs = sched.Costs(), p2p: str = "",
                 executor: str = "", max_batch: int = 0):
        self.stage, self.topo, self.rank, self.groups = stage, topo, rank, groups
        self.pp_idx, self.dp_idx, self.tp_idx = topo.coords(rank)
        self.p = topo.pp
        self.m = microbatches
        self.schedule = schedule
        self.timeout_s = timeout_s
        self.costs = costs
        self.prev_rank = topo.prev_stage_rank(rank)
        self.next_rank = topo.next_stage_rank(rank)
        self._ops_cache: Dict[int, List[Op]] = {}
        self._send_works: List[_Pending] = []
        self.is_cuda = stage.device.type == "cuda"
        self.steps_done = 0
        self._in_step = False
        self.lookahead = int(os.environ.get("SDML_LOOKAHEAD", lookahead))
        self.tracer = None       # utils.trace.StepTracer; attach to get a per-op timeline
        self._fb_streams = None  # (forward, backward) streams of latency-bound stages, see train_step
        self._fb_events: Dict[object, torch.cuda.Event] = {}
        # SDML_P2P=native: stage-boundary transfers through csrc/p2p_engine.cu (own NCCL
        # communicators + side streams, event-ordered) instead of torch.distributed's NCCL backend
        self.native = None
        self.bnd = None          # parallel.boundary.SymmBoundary when p2p_mode == "symm"
        self.p2p_mode = p2p or os.environ.get("SDML_P2P", "auto")
        multi = self.groups is not None and self.p > 1
        symm_ok = (multi and self.is_cuda and topo.tp == 1 and max_batch > 0
                   and getattr(stage, "supports_symm_boundary", False))
        if self.p2p_mode == "auto":
            # GPUs: push into peer-mapped slots; everything else: torch.distributed p2p
            self.p2p_mode = "symm" if symm_ok else "torch"
        elif self.p2p_mode == "symm" and multi and not symm_ok:
            raise ValueError("--p2p symm needs CUDA, tp == 1 and a stage with symmetric receive slots "
                             f"({type(stage).__name__}, tp={topo.tp})")
        # executor: "python" (this class walks the op list), "native" (csrc/executor.cu; on CUDA the
        # Trainer switches to per-op CUDA graphs, elsewhere the C++ loop calls back into this
        # class's helpers through hooks)
        self.executor = executor or os.environ.get("SDML_EXECUTOR", "python")
        if multi and self.p2p_mode == "symm":
            from .boundary import SymmBoundary
            self.bnd = SymmBoundary(topo, rank, self.groups.ctrl, stage.device, stage, max_batch, timeout_s)
            stage.boundary = self.bnd
        if (multi and self.is_cuda and self.p2p_mode == "native"):
            from .p2p_native import PipelineChannels
            self.native = PipelineChannels(topo, rank, self.groups.ctrl, stage.device)
        if multi and self.native is None and self.bnd is None:
            self._warmup_p2p()

    def _warmup_p2p(self) -> None:
        """Establish every p2p connection once, in one global order.

        NCCL sets up the transport between two ranks lazily, inside the FIRST send/recv on a
        communicator, and that set-up blocks the host until the peer shows up.  With receives
        posted ahead of time a rank's first touch of the gradient communicator can come before
        its first touch of the activation communicator while the neighbour does the opposite:
        both hosts then sit in a different set-up forever.  Walking the boundaries in the same
        order on every rank (activation direction, then gradient direction) removes that."""
        dev = self.stage.device
        token = torch.zeros(8, dtype=torch.float32, device=dev)
        for b in range(self.p - 1):
            if self.pp_idx == b:
                dist.send(token, dst=self.next_rank, group=self.groups.fwd_p2p[b])
                dist.recv(token, src=self.next_rank, group=self.groups.bwd_p2p[b])
            elif self.pp_idx == b + 1:
                dist.recv(token, src=self.prev_rank, group=self.groups.fwd_p2p[b])
                dist.send(token, dst=self.prev_rank, group=self.groups.bwd_p2p[b])
            if self.is_cuda:
                torch.cuda.synchronize(dev)

    def _fb_event(self, key) -> "torch.cuda.Event":
        ev = self._fb_events.get(key)
        if ev is None:
            ev = self._fb_events[key] = torch.cuda.Event()
        return ev

    # ------------------------------------------------------------------ schedule
    def ops_for(self, n_mb: int) -> List[Op]:
        ops = self._ops_cache.get(n_mb)
        if ops is None:
            kind = self.schedule
            if kind == "lockstep" and n_mb != 1:
                kind = "gpipe"
            ops = sched.generate(kind, self.p, n_mb, self.pp_idx, self.costs)
            sched.validate(ops, n_mb)
            self._ops_cache[n_mb] = ops
        return ops

    # ------------------------------------------------------------------ p2p helpers
    def _wait(self, pend: _Pending) -> None:
        if self.bnd is not None:
            if pend.work is not None:      # receives only; a send has nothing left to wait for
                self.bnd.wait(*pend.work)  # the current stream waits for the slot's flag
            return
        if self.native is not None:
            chan, ticket = pend.work
            chan.wait(ticket)              # the current stream waits for the transfer's event
            return
        if self.is_cuda:
            # NCCL: wait() makes the current stream wait; the host does not block.  A dead peer
            # surfaces through the process group's watchdog, whose timeout init_distributed sets
            # to --timeout_s (train.py); the symmetric boundary and the fused collectives carry
            # their own in-kernel deadlines (Trainer.check_peers).
            pend.work.wait()
            return
        # gloo: is_completed() of a receive never turns true before wait(), so bound the
        # blocking wait itself.
        try:
            pend.work.wait(datetime.timedelta(seconds=self.timeout_s))
        except RuntimeError as e:
            raise PipelineTimeout(
                f"rank {self.rank}: gave up after {self.timeout_s:.0f}s waiting for {pend.what}: {e}") from e

    def _post_recv_fwd(self, mb: int, size: int) -> _Pending:
        if self.bnd is not None:
            # nothing to post: the previous stage writes straight into this rank's slot
            return _Pending(("act", mb), self.stage.input_buffer(mb), f"activation of micro-batch {mb}")
        buf = self.stage.input_buffer(mb) if self._in_step else None
        if buf is None:
            buf = torch.empty(self.stage.in_shape(size), dtype=self.stage.act_dtype, device=self.stage.device)
        what = f"activation of micro-batch {mb} from rank {self.prev_rank}"
        if self.native is not None:
            ch = self.native.act_in
            return _Pending((ch, ch.recv(buf, 0)), buf, what)
        g = self.groups.fwd_p2p[self.pp_idx - 1]
        w = dist.irecv(buf, src=self.prev_rank, group=g)
        return _Pending(w, buf, what)

    def _post_recv_bwd(self, mb: int, size: int) -> _Pending:
        if self.bnd is not None:
            return _Pending(("grad", mb), self.stage.grad_buffer(mb), f"gradient of micro-batch {mb}")
        buf = self.stage.grad_buffer(mb) if self._in_step else None
        if buf is None:
            buf = torch.empty(self.stage.out_shape(size), dtype=self.stage.act_dtype, device=self.stage.device)
        what = f"gradient of micro-batch {mb} from rank {self.next_rank}"
        if self.native is not None:
            ch = self.native.grad_in
            return _Pending((ch, ch.recv(buf, 0)), buf, what)
        g = self.groups.bwd_p2p[self.pp_idx]
        w = dist.irecv(buf, src=self.next_rank, group=g)
        return _Pending(w, buf, what)

    def _send_fwd(self, mb: int, y: torch.Tensor) -> None:
        what = f"send of activation {mb} to rank {self.next_rank}"
        if self.bnd is not None:
            self.bnd.send("act", mb, y)
            return
        y = y.contiguous()
        if self.native is not None:
            ch = self.native.act_out       # sender is rank 0 of its channel, receiver rank 1
            self._send_works.append(_Pending((ch, ch.send(y, 1)), y, what))
            return
        g = self.groups.fwd_p2p[self.pp_idx]
        w = dist.isend(y, dst=self.next_rank, group=g)
        self._send_works.append(_Pending(w, y, what))

    def _send_bwd(self, mb: int, dx: torch.Tensor) -> None:
        what = f"send of gradient {mb} to rank {self.prev_rank}"
        if self.bnd is not None:
            self.bnd.send("grad", mb, dx)
            return
        dx = dx.contiguous()
        if self.native is not None:
            ch = self.native.grad_out
            self._send_works.append(_Pending((ch, ch.send(dx, 1)), dx, what))
            return
        g = self.groups.bwd_p2p[self.pp_idx - 1]
        w = dist.isend(dx, dst=self.prev_rank, group=g)
        self._send_works.append(_Pending(w, dx, what))

    def _drain_sends(self) -> None:
        for pend in self._send_works:
            self._wait(pend)
        self._send_works.clear()

    # ------------------------------------------------------------------ training step
    def train_step(self, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor],
                   batch_size: int) -> Optional[torch.Tensor]:
        """One optimizer step over ``batch_size`` samples of this DP replica.

        ``inputs`` is needed on the first stage, ``targets`` on the last; both are already
        on the stage's device.  Returns the mean loss as a device tensor on the last stage
        (no host sync), ``None`` elsewhere.
        """
        stage = self.stage
        sizes = split_microbatches(batch_size, self.m)
        n_mb = len(sizes)
        offs = [0]
        for s in sizes:
            offs.append(offs[-1] + s)
        ops = self.ops_for(n_mb)
        if self.executor == "native" and self.tracer is None:
            return self._train_step_native_loop(ops, inputs, targets, batch_size, sizes, offs)
        tracer = self.tracer
        span = tracer.span if tracer is not None else _no_span
        if tracer is not None:
            tracer.begin_step()
        if self.bnd is not None:
            self.bnd.begin_round(sizes)
        stage.begin_step(batch_size, sizes)
        self._in_step = True

        needs_recv = []
        for i, op in enumerate(ops):
            if op.kind == OpKind.F and not stage.is_first:
                needs_recv.append(i)
            elif op.kind in (OpKind.B, OpKind.BW) and not stage.is_last:
                needs_recv.append(i)
        posted: Dict[int, _Pending] = {}
        next_recv = 0

        def post_upto(i_op: int) -> None:
            nonlocal next_recv
            while next_recv < len(needs_recv) and needs_recv[next_recv] <= i_op + self.lookahead:
                j = needs_recv[next_recv]
                o = ops[j]
                if o.kind == OpKind.F:
                    posted[j] = self._post_recv_fwd(o.mb, sizes[o.mb])
                else:
                    posted[j] = self._post_recv_bwd(o.mb, sizes[o.mb])
                next_recv += 1

        # Latency-bound stages (a handful of tiny kernels per op, e.g. the reference model at 7 rows
        # per micro-batch) leave the GPU almost empty: their forward and backward ops are issued on
        # TWO streams so that F(i+1) overlaps B(i).  Order is kept by events only where the data
        # needs it (B(i) after F(i)); everything joins the main stream before the optimizer.
        two = (self.bnd is not None and tracer is None and n_mb > 1
               and getattr(stage, "concurrent_fb", False))
        if two:
            if self._fb_streams is None:
                self._fb_streams = (torch.cuda.Stream(device=stage.device), torch.cuda.Stream(device=stage.device))
                self._fb_events = {}
            sF, sB = self._fb_streams
            main = torch.cuda.current_stream(stage.device)
            fork = self._fb_event("fork")
            fork.record(main)
            sF.wait_event(fork)
            sB.wait_event(fork)
            ctxF, ctxB = torch.cuda.stream(sF), torch.cuda.stream(sB)
        for i, op in enumerate(ops):
            post_upto(i)
            mb = op.mb
            if op.kind == OpKind.F:
                with (ctxF if two else contextlib.nullcontext()):
                    if stage.is_first:
                        x = inputs[offs[mb]:offs[mb + 1]]
                    else:
                        pend = posted.pop(i)
                        self._wait(pend)
                        x = pend.buf
                    t = targets[offs[mb]:offs[mb + 1]] if stage.is_last else None
                    with span("F", mb):
                        y = stage.forward(mb, x, t)
                    if not stage.is_last:
                        self._send_fwd(mb, y)
                    if two:
                        self._fb_event(("F", mb)).record(sF)
            elif op.kind in (OpKind.B, OpKind.BW):
                with (ctxB if two else contextlib.nullcontext()):
                    if two:
                        sB.wait_event(self._fb_event(("F", mb)))
                    if stage.is_last:
                        dy = None
                    else:
                        pend = posted.pop(i)
                        self._wait(pend)
                        dy = pend.buf
                    with span("B", mb):
                        dx = stage.backward_input(mb, dy)
                    if not stage.is_first:
                        self._send_bwd(mb, dx)
                    if op.kind == OpKind.BW:
                        with span("W", mb):
                            stage.backward_weight(mb)
            elif op.kind == OpKind.W:
                with (ctxB if two else contextlib.nullcontext()):
                    with span("W", mb):
                        stage.backward_weight(mb)
        if two:
            main.wait_stream(sF)
            main.wait_stream(sB)
        with span("opt"):
            stage.end_step()
        if self.bnd is not None:
            self.bnd.end_round()       # every reader of this step's slots is enqueued: acknowledge
        self._drain_sends()
        if tracer is not None:
            tracer.end_step()
        self._in_step = False
        self.steps_done += 1
        return stage.mean_loss() if stage.is_last else None

    def _train_step_native_loop(self, ops, inputs, targets, batch_size, sizes, offs):
        """Same step, but the op list is walked by the C++ loop (``sdml_exec_run_hooks``) and this
        class only supplies the actions.  On GPUs the Trainer replaces this by the fully native
        plan (per-op CUDA graphs) after the first steps; this path is what runs on CPUs and while
        warming up, and it keeps the two executors on one set of helpers."""
        from .native_exec import run_with_hooks
        stage = self.stage
        if self.bnd is not None:
            self.bnd.begin_round(sizes)
        stage.begin_step(batch_size, sizes)
        self._in_step = True
        recvs: Dict[int, _Pending] = {}
        sends: Dict[int, _Pending] = {}
        acts: Dict[int, torch.Tensor] = {}
        grads: Dict[int, torch.Tensor] = {}
        counter = [0]

        def ticket(table, pend):
            counter[0] += 1
            table[counter[0]] = pend
            return counter[0]

        def post_recv(direction, mb):
            pend = (self._post_recv_fwd if direction == 0 else self._post_recv_bwd)(mb, sizes[mb])
            return ticket(recvs, pend)

        def wait_recv(direction, mb, t):
            pend = recvs.pop(t)
            self._wait(pend)
            (acts if direction == 0 else grads)[mb] = pend.buf

        def compute(kind, mb):
            if kind == 0:
                x = inputs[offs[mb]:offs[mb + 1]] if stage.is_first else acts.pop(mb)
                tgt = targets[offs[mb]:offs[mb + 1]] if stage.is_last else None
                acts[("out", mb)] = stage.forward(mb, x, tgt)
            elif kind == 1:
                grads[("out", mb)] = stage.backward_input(mb, None if stage.is_last else grads.pop(mb))
            else:
                stage.backward_weight(mb)

        def send(direction, mb):
            before = len(self._send_works)
            if direction == 0:
                self._send_fwd(mb, acts.pop(("out", mb)))
            else:
                self._send_bwd(mb, grads.pop(("out", mb)))
            if self.bnd is not None:       # pushed on the compute stream: nothing to wait for
                return ticket(sends, _Pending(None, None, "pushed"))
            return ticket(sends, self._send_works[before])

        def wait_send(direction, mb, t):
            self._wait(sends.pop(t))

        def end_step():
            stage.end_step()
            if self.bnd is not None:
                self.bnd.end_round()

        run_with_hooks(ops, stage.is_first, stage.is_last, self.lookahead, compute=compute,
                       post_recv=post_recv, wait_recv=wait_recv, send=send, wait_send=wait_send,
                       end_step=end_step)
        self._send_works.clear()
        self._in_step = False
        self.steps_done += 1
        return stage.mean_loss() if stage.is_last else None

    # ------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def eval_batch(self, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor],
                   batch_size: int) -> None:
        """Forward-only pass of one batch; loss-sum / correct-count accumulate on the last
        stage (read them once per epoch)."""
        stage = self.stage
        if self.bnd is not None:
            x = inputs if stage.is_first else self.bnd.eval_recv(batch_size, stage.in_shape(1)[1:])
            y = stage.eval_forward(x, targets if stage.is_last else None)
            if not stage.is_last:
                self.bnd.eval_send(y.contiguous())
            self.bnd.eval_end()
            return
        if stage.is_first:
            x = inputs
        else:
            pend = self._post_recv_fwd(0, batch_size)
            self._wait(pend)
            x = pend.buf
        y = stage.eval_forward(x, targets if stage.is_last else None)
        if not stage.is_last:
            self._send_fwd(0, y)
            self._drain_sends()
