"""Training / evaluation driver: the public entry point of the framework.

``run(cfg)`` is what ``simple_distributed.py --rank R ...`` executes on every rank.  It
covers what the reference's ``__main__`` + ``run_master`` do
(/root/reference/simple_distributed.py:86-136, 138-186): rendezvous from
``--master_addr/--master_port`` on ``--interface``, build this rank's stage, iterate
epochs of train + test with the reference's log lines -- but as SPMD: every rank builds
its own stage and runs the same loop (the reference's rank 1 only serves RPCs), the loss
lives on the last stage, and the optimizer is local to each rank.
"""
from __future__ import annotations

import datetime
import os
import sys
import time
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from .models.lenet import LeNetStage0, LeNetStage1, build_lenet_stage
from .models.mlp import MLPStage
from .parallel.pipeline import PipelineEngine
from .parallel.stage import AutogradStage, Stage
from .parallel.topology import ProcessGroups, Topology, build_process_groups
from .utils import checkpoint as ckpt
from .utils import logging as slog
from .utils.config import TrainConfig, export_env
from .utils.data import ArrayDataset, BatchLoader, load_mnist_like, synthetic_mlp


def pick_device(cfg: TrainConfig) -> torch.device:
    want = cfg.device
    if want == "auto":
        want = "cuda" if torch.cuda.is_available() else "cpu"
    if want == "cuda":
        local = int(os.environ.get("LOCAL_RANK", cfg.rank % max(1, torch.cuda.device_count())))
        torch.cuda.set_device(local)
        if cfg.dtype == "fp32":
            # the reference computes in fp32 on the CPU; keep cuDNN / cuBLAS library calls of
            # the conv stage in true fp32 rather than TF32
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False
        return torch.device("cuda", local)
    return torch.device("cpu")


def pick_backend(cfg: TrainConfig, device: torch.device) -> str:
    if cfg.backend != "auto":
        return cfg.backend
    return "nccl" if device.type == "cuda" else "gloo"


def init_distributed(cfg: TrainConfig, device: torch.device) -> Tuple[Topology, Optional[ProcessGroups]]:
    topo = Topology(cfg.world_size, cfg.resolved_pp(), cfg.dp, cfg.tp)
    if cfg.world_size == 1:
        return topo, None
    export_env(cfg)
    backend = pick_backend(cfg, device)
    if not dist.is_initialized():
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = device
        dist.init_process_group(backend=backend, rank=cfg.rank, world_size=cfg.world_size,
                                timeout=datetime.timedelta(seconds=max(30.0, cfg.timeout_s)), **kwargs)
    groups = build_process_groups(topo, cfg.rank, backend)
    return topo, groups


def build_stage(cfg: TrainConfig, topo: Topology, rank: int, device: torch.device) -> Stage:
    pp_idx, dp_idx, tp_idx = topo.coords(rank)
    pp = topo.pp
    if cfg.model == "lenet" and cfg.fused_lenet and device.type == "cuda":
        from .models.lenet_fused import LeNetFusedStage
        return LeNetFusedStage(pp_idx, pp, device, dropout=cfg.dropout, ref_eval_dropout=cfg.ref_eval_dropout,
                               seed=cfg.seed, optimizer=cfg.optimizer, lr=cfg.lr, momentum=cfg.momentum,
                               weight_decay=cfg.weight_decay, replica=dp_idx)
    if cfg.model == "lenet":
        module = build_lenet_stage(pp_idx, pp, dropout=cfg.dropout,
                                   ref_eval_dropout=cfg.ref_eval_dropout, seed=cfg.seed, replica=dp_idx)
        if pp == 1:
            in_f, out_f, prefixes = (1, 28, 28), (10,), {0: "stage0.", 1: "stage1."}
        elif pp_idx == 0:
            in_f, out_f, prefixes = (1, 28, 28), (LeNetStage0.out_features,), None
        else:
            in_f, out_f, prefixes = (LeNetStage1.in_features,), (LeNetStage1.num_classes,), None
        return AutogradStage(module, pp_idx == 0, pp_idx == pp - 1, device, in_f, out_f,
                             optimizer=cfg.optimizer, lr=cfg.lr, momentum=cfg.momentum,
                             weight_decay=cfg.weight_decay, part_prefixes=prefixes)
    if cfg.model == "mlp" and cfg.virtual_stages > 1:
        from .parallel.interleaved import ChunkedStage
        v = cfg.virtual_stages
        chunks = [MLPStage(c * pp + pp_idx, pp * v, cfg.layers, cfg.hidden, cfg.classes, device,
                           optimizer=cfg.optimizer, lr=cfg.lr, momentum=cfg.momentum,
                           weight_decay=cfg.weight_decay, seed=cfg.seed, dtype=cfg.dtype) for c in range(v)]
        return ChunkedStage(chunks, pp_idx, pp)
    if cfg.model == "mlp":
        if topo.tp > 1:
            from .parallel.tensor_parallel import TPMLPStage
            return TPMLPStage(pp_idx, pp, cfg.layers, cfg.hidden, cfg.classes, device,
                              optimizer=cfg.optimizer, lr=cfg.lr, momentum=cfg.momentum,
                              weight_decay=cfg.weight_decay, seed=cfg.seed, tp=topo.tp, tp_idx=tp_idx,
                              dtype=cfg.dtype, backend="fused" if cfg.fused_collectives else "nccl")
        return MLPStage(pp_idx, pp, cfg.layers, cfg.hidden, cfg.classes, device,
                        optimizer=cfg.optimizer, lr=cfg.lr, momentum=cfg.momentum,
                        weight_decay=cfg.weight_decay, seed=cfg.seed, dtype=cfg.dtype)
    raise ValueError(cfg.model)


def build_data(cfg: TrainConfig) -> Tuple[ArrayDataset, ArrayDataset]:
    if cfg.model == "lenet":
        train = load_mnist_like(cfg.data_root, True, cfg.train_samples, cfg.synthetic, cfg.seed)
        test = load_mnist_like(cfg.data_root, False, cfg.test_samples, cfg.synthetic, cfg.seed)
    else:
        classes = cfg.classes or cfg.hidden
        train = synthetic_mlp(cfg.train_samples, cfg.hidden, classes, seed=cfg.seed)
        test = synthetic_mlp(cfg.test_samples, cfg.hidden, classes, seed=cfg.seed + 1)
    return train, test


class LossHandle:
    """The loss of one step on its way to the host: an asynchronous device->host copy into pinned
    memory plus the event that marks its arrival.  ``float(h)`` / ``h.item()`` blocks only until
    THAT step has finished, so a training loop can read step k's loss after it has launched step
    k+1 and keep the GPU busy (the reference reads ``loss[0]`` synchronously, ref:114-117)."""
    __slots__ = ("_buf", "_ev", "_val")

    def __init__(self, buf: torch.Tensor, ev):
        self._buf, self._ev, self._val = buf, ev, None

    def item(self) -> float:
        if self._val is None:
            if self._ev is not None:
                self._ev.synchronize()
            self._val = float(self._buf)
        return self._val

    __float__ = item


class Trainer:
    """Everything one rank needs; also used directly by the tests and the benchmark."""

    def __init__(self, cfg: TrainConfig, out=sys.stdout):
        self.cfg = cfg.validate()
        self.out = out
        self.device = pick_device(cfg)
        self.topo, self.groups = init_distributed(cfg, self.device)
        self.rank = cfg.rank
        self.pp_idx, self.dp_idx, self.tp_idx = self.topo.coords(self.rank)
        if self.device.type == "cuda" and cfg.world_size > 1:
            from .ops import _lib as _native
            _native.set_wait_timeout(cfg.timeout_s)     # deadline of every in-kernel wait for a peer GPU
        self.stage = build_stage(cfg, self.topo, self.rank, self.device)
        self.stage.timeout_s = cfg.timeout_s      # bound of the in-kernel waits of fused collectives
        if self.groups is not None and self.topo.tp > 1 and hasattr(self.stage, "attach_tp"):
            self.stage.attach_tp(self.groups.tp, self.topo.tp_group_ranks(self.rank), self.rank,
                                 self.groups.ctrl)
        if self.groups is not None and self.topo.dp > 1:
            self.stage.dp_group, self.stage.dp_size = self.groups.dp, self.topo.dp
            if cfg.fused_collectives and self.device.type == "cuda" and hasattr(self.stage, "enable_fused_dp"):
                self.stage.enable_fused_dp(self.groups, self.topo, self.rank)
        if cfg.virtual_stages > 1:
            from .parallel.interleaved import InterleavedEngine
            self.engine = InterleavedEngine(self.stage, self.topo, self.rank, cfg.microbatches,
                                            timeout_s=cfg.timeout_s,
                                            backend=pick_backend(cfg, self.device) if self.groups is not None else None)
        else:
            self.engine = self._make_engine(cfg)
        self.tracer = None
        self._finish_init(cfg)

    def _make_engine(self, cfg: TrainConfig) -> PipelineEngine:
        # executor "auto" / "graph": the engine itself runs the Python loop (that is what gets captured)
        return PipelineEngine(self.stage, self.topo, self.rank, self.groups,
                                     microbatches=cfg.microbatches, schedule=cfg.schedule,
                                     timeout_s=cfg.timeout_s,
                                     p2p="native" if (cfg.executor == "native" and cfg.p2p == "auto") else cfg.p2p,
                                     executor="native" if cfg.executor == "native" else "python",
                                     max_batch=cfg.batch_size)

    def _finish_init(self, cfg: TrainConfig) -> None:
        if cfg.trace_dir:
            from .utils.trace import StepTracer
            # the first steps carry lazy allocations / connection set-up: keep them out of the timeline
            self.tracer = StepTracer(self.device, self.rank, self.pp_idx, skip=3, max_steps=cfg.trace_steps)
            self.engine.tracer = self.tracer
        self.train_ds, self.test_ds = build_data(cfg)
        gb = cfg.batch_size * self.topo.dp
        self.train_loader = BatchLoader(self.train_ds, gb)
        self.test_loader = BatchLoader(self.test_ds, cfg.batch_size)
        self.owns_loss = self.stage.is_last and self.dp_idx == 0 and self.tp_idx == 0
        self.metrics = slog.JsonlSink(cfg.metrics_jsonl) if (cfg.metrics_jsonl and self.owns_loss) else None
        self.start_epoch = 1
        self.global_step = 0
        self.samples_seen = 0
        self.train_time_s = 0.0
        # CUDA graphs: the launch-bound single-GPU step (the reference model at batch 60 is
        # ~50 tiny kernels) is captured once per batch size and replayed.  Adam's bias
        # correction is a host-side scalar, so only SGD steps are captured.
        self._copy_stream = None
        self._throttle_evs = None
        self._throttle_i = 0
        self._peer_poll = 0
        self._graphs: Dict[int, tuple] = {}
        # second set of static graph inputs per batch size: the H2D copy of step k+1 (copy stream)
        # overlaps the replay of step k.
        self._graphs_alt: Dict[int, tuple] = {}
        # (with DP replicas the second capture's warm-up steps are collective: every replica reaches it
        # at the same call of train_batch, so they stay in lock-step)
        # On pipelines only the first stage has a batch to copy (the others replay their one graph); its
        # second graph is a late capture, which executes nothing and involves no other rank.
        self._double_buffer = (self.topo.tp == 1 and (self.topo.pp == 1 or self.stage.is_first)
                               and os.environ.get("SDML_E2E_DOUBLE_BUFFER", "1") == "1")
        self._parity = 0
        self._db_events: Dict[int, object] = {}
        self._db_prev_launch = None
        self._db_markers = None
        self._db_marker_idx = 0
        # SDML_EXECUTOR=native: after three eager steps the rank's op list is captured op by op
        # into CUDA graphs and every later step is one call into the C++ executor
        self._native_exec = self.engine.executor == "native" and self.device.type == "cuda"
        if self._native_exec and hasattr(self.stage, "static_slots"):
            self.stage.static_slots = True      # per-op CUDA graphs need fixed receive addresses
        self._runners: Dict[int, object] = {}
        self._loss_ring = None      # pinned scalars + events for asynchronous loss read-back
        self._loss_slot = 0
        # With DP replicas the captured step ends in the fused all-reduce + SGD kernel (its epoch
        # is device-resident); pipeline p2p is not captured.
        # A pipeline rank's step is capturable too when its transfers are pushes into peer-mapped slots
        # (p2p "symm": only stream-ordered kernels, round counters on the device).
        dp_ok = self.topo.dp == 1 or getattr(self.stage, "_fused_dp", None) is not None
        pp_ok = self.topo.pp == 1 or self.engine.bnd is not None
        stage_ok = hasattr(self.stage, "module") or (isinstance(self.stage, MLPStage)
                                                     and getattr(self.stage, "_dp_w", None) is None)
        self.graph_ok = (cfg.use_cuda_graph and cfg.executor in ("auto", "graph")
                         and self.device.type == "cuda" and pp_ok
                         and self.topo.tp == 1 and dp_ok and cfg.optimizer == "sgd"
                         and stage_ok and self.tracer is None)
        if cfg.executor == "graph" and not self.graph_ok:
            raise ValueError("--executor graph: this configuration cannot be captured (needs CUDA, SGD, tp == 1, "
                             "no tracing, --p2p symm for pp > 1, and DP through the in-kernel all-reduce)")
        # Late capture: the first three steps of a batch size run eagerly (they are the warm-up CUDA
        # graphs need), the fourth call is captured without executing -- no snapshot / restore of the
        # training state, nothing collective about it.  Used for pipelines and for stages too big to
        # snapshot; the single-GPU reference model keeps the immediate capture (state-neutral warm-up).
        self._late_capture = self.topo.pp > 1 or not hasattr(self.stage, "module")
        self._eager_steps: Dict[int, int] = {}
        self.launches_per_graph: Dict[int, int] = {}
        if cfg.resume and cfg.ckpt_dir:
            self.load_checkpoint()

    # ---- data movement -----------------------------------------------------------------
    def _replica_slice(self, data: torch.Tensor, target: torch.Tensor):
        """This DP replica's rows of a global batch (ragged tail split as evenly as possible)."""
        n, dp = data.shape[0], self.topo.dp
        if dp == 1:
            return data, target, n
        base, rem = divmod(n, dp)
        lo = self.dp_idx * base + min(self.dp_idx, rem)
        hi = lo + base + (1 if self.dp_idx < rem else 0)
        return data[lo:hi], target[lo:hi], hi - lo

    def _tp_rows(self, t: torch.Tensor) -> torch.Tensor:
        """Rows of a replica batch that this TP rank feeds / scores: within every micro-batch
        the rows are split into ``tp`` contiguous shards (the activation is row-sharded between
        the fused all-gather->GEMM / GEMM->reduce-scatter pairs)."""
        tp = self.topo.tp
        if tp == 1:
            return t
        m = self.cfg.microbatches
        n = t.shape[0]
        assert n % (m * tp) == 0, "batch must split evenly into micro-batches x TP shards"
        per = n // m
        sh = per // tp
        # one (asynchronous, if `t` is pinned) H2D copy per micro-batch shard, no host-side gather
        out = torch.empty((m * sh, *t.shape[1:]), dtype=t.dtype, device=self.device)
        for i in range(m):
            lo = i * per + self.tp_idx * sh
            out[i * sh:(i + 1) * sh].copy_(t[lo:lo + sh], non_blocking=True)
        return out

    def _h2d(self, t: torch.Tensor) -> torch.Tensor:
        """Host -> device on a dedicated copy stream, so the transfer of the next step's batch
        overlaps whatever the compute stream still has queued; the compute stream then waits
        for exactly this copy."""
        if self.device.type != "cuda":
            return t.to(self.device)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        with torch.cuda.stream(self._copy_stream):
            out = t.to(self.device, non_blocking=True)
        torch.cuda.current_stream(self.device).wait_stream(self._copy_stream)
        out.record_stream(torch.cuda.current_stream(self.device))
        return out

    def _throttle(self) -> None:
        """Let the host run at most two steps ahead of the device (bounds queued input copies)."""
        if self.device.type != "cuda":
            return
        if self._throttle_evs is None:
            self._throttle_evs = [torch.cuda.Event() for _ in range(3)]
        ev = self._throttle_evs[self._throttle_i % 3]       # last recorded three calls ago
        if self._throttle_i >= 3:
            ev.synchronize()
            self.check_peers()
        ev.record(torch.cuda.current_stream(self.device))
        self._throttle_i += 1

    def check_peers(self) -> None:
        """Raise ``PipelineTimeout`` if a device-side wait for a pipeline neighbour has given up."""
        from .parallel.pipeline import PipelineTimeout
        if self.device.type == "cuda":
            from .ops import _lib as _native
            try:
                _native.check_waits()
            except _native.NativeError as e:
                raise PipelineTimeout(f"rank {self.rank}: {e}") from e
        if hasattr(self.stage, "check_peers"):
            try:
                self.stage.check_peers()
            except RuntimeError as e:
                raise PipelineTimeout(f"rank {self.rank}: {e}") from e
        for name in ("_fused_dp", "_dp_b"):
            fd = getattr(self.stage, name, None)
            if fd is not None and hasattr(fd, "check"):
                try:
                    fd.check()
                except RuntimeError as e:
                    raise PipelineTimeout(f"rank {self.rank}: {e}") from e
        bnd = self.engine.bnd
        if bnd is not None:
            from .parallel.boundary import BoundaryTimeout
            try:
                bnd.check()
            except BoundaryTimeout as e:
                raise PipelineTimeout(str(e)) from e

    def _to_device(self, data, target, row_shard: bool = True):
        if self.topo.tp > 1 and row_shard:
            x = self._tp_rows(data) if self.stage.is_first else None
            t = self._tp_rows(target) if self.stage.is_last else None
            return x, t
        x = self._h2d(data) if self.stage.is_first else None
        t = self._h2d(target) if self.stage.is_last else None
        return x, t

    # ---- one optimizer step -------------------------------------------------------------
    def _capture(self, n: int, x: Optional[torch.Tensor], t: Optional[torch.Tensor]):
        """Capture ``engine.train_step`` for batch size ``n`` into a CUDA graph (static input
        buffers; three eager warm-up steps on a side stream first, as CUDA graphs require)."""
        from .ops import _lib
        xs = torch.empty_like(x) if x is not None else None
        ts = torch.empty_like(t) if t is not None else None
        if xs is not None:
            xs.copy_(x)
        if ts is not None:
            ts.copy_(t)
        stage = self.stage
        opt = stage.optimizer
        if self._late_capture:
            # the eager steps before this call were the warm-up; capturing executes nothing, so only
            # the host-side counters have to be put back
            snap_steps = (opt.step_count, self.engine.steps_done)
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            before = _lib.launch_count()
            with torch.cuda.graph(graph):
                loss = self.engine.train_step(xs, ts, n)
            self.launches_per_graph[n] = _lib.launch_count() - before
            opt.step_count, self.engine.steps_done = snap_steps
            entry = (graph, xs, ts, loss)
            self._graphs[n] = entry
            return entry
        # warm-up + capture must not change the training state: snapshot, then restore
        snap_p = [p.detach().clone() for p in stage.params]
        rngs = {}
        for name in ("rng_step", "rng_s0", "rng_s1"):      # device-resident dropout counters
            r = getattr(stage, name, None)
            if r is not None:
                rngs[r.data_ptr()] = r
        snap_rng = [(r, r.clone()) for r in rngs.values()]
        snap_m = [b.clone() for b in getattr(opt, "momentum_buffers", [])]
        snap_buf = [b.clone() for b in stage.module.buffers()]
        snap_steps = (opt.step_count, self.engine.steps_done)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):
                self.engine.train_step(xs, ts, n)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        before = _lib.launch_count()
        with torch.cuda.graph(graph):
            loss = self.engine.train_step(xs, ts, n)
        self.launches_per_graph[n] = _lib.launch_count() - before
        with torch.no_grad():
            for p, sp in zip(stage.params, snap_p):
                p.copy_(sp)
            for b, sb in zip(getattr(opt, "momentum_buffers", []), snap_m):
                b.copy_(sb)
            for b, sb in zip(stage.module.buffers(), snap_buf):
                b.copy_(sb)
            for r, saved in snap_rng:
                r.copy_(saved)
        opt.step_count, self.engine.steps_done = snap_steps
        entry = (graph, xs, ts, loss)
        self._graphs[n] = entry
        return entry

    def train_step_device(self, x: Optional[torch.Tensor], t: Optional[torch.Tensor], n: int):
        """One optimizer step from tensors already on the device; returns the mean loss as a
        device tensor on the loss-owning rank (no host sync)."""
        if not self.graph_ok:
            if self._native_exec:
                runner = self._runners.get(n)
                if runner is None and self.engine.steps_done >= 3:
                    runner = self._build_runner(x, t, n)
                if runner is not None:
                    return runner.step(x, t)
            return self.engine.train_step(x, t, n)
        entry = self._graphs.get(n)
        if entry is None:
            if self._late_capture and self._eager_steps.get(n, 0) < 3:
                self._eager_steps[n] = self._eager_steps.get(n, 0) + 1
                return self.engine.train_step(x, t, n)
            entry = self._capture(n, x, t)     # leaves the training state untouched
        graph, xs, ts, loss = entry
        if xs is not None:
            xs.copy_(x, non_blocking=True)
        if ts is not None:
            ts.copy_(t, non_blocking=True)
        graph.replay()
        self.stage.optimizer.step_count += 1
        self.engine.steps_done += 1
        self._db_prev_launch = None     # (see _replay_double_buffered)
        return loss

    def _build_runner(self, x, t, n: int):
        from .parallel.native_exec import NativeStepRunner
        why = NativeStepRunner.supported(self.engine)
        if why is None and getattr(self.stage, "_dp_w", None) is not None:
            why = "fused DP weight-gradient path is not captured"
        if why is not None:
            print(f"[native executor] rank {self.rank}: staying on the Python executor: {why}", file=sys.stderr)
            self._native_exec = False
            return None
        runner = NativeStepRunner(self.engine, x, t, n)
        self._runners[n] = runner
        return runner

    def _async_loss(self, loss: Optional[torch.Tensor]) -> Optional[LossHandle]:
        """Queue the D2H copy of ``loss`` (a device scalar that the next step may overwrite) into
        a pinned slot and return the handle that waits for it."""
        if loss is None:
            return None
        if self.device.type != "cuda":
            return LossHandle(loss.detach().clone(), None)
        if self._loss_ring is None:
            k = 8
            self._loss_ring = (torch.zeros(k, dtype=torch.float32).pin_memory(),
                               [torch.cuda.Event() for _ in range(k)])
        bufs, evs = self._loss_ring
        i = self._loss_slot
        self._loss_slot = (i + 1) % len(evs)
        bufs[i:i + 1].copy_(loss.reshape(1), non_blocking=True)
        evs[i].record(torch.cuda.current_stream(self.device))
        return LossHandle(bufs[i:i + 1], evs[i])

    def _replay_double_buffered(self, n: int, d: torch.Tensor, t: torch.Tensor):
        """Graph replay with alternating static input sets: this step's H2D copies run on the copy
        stream while the previous step's graph is still executing."""
        if n not in self._graphs_alt:
            primary = self._graphs[n]
            self._graphs_alt[n] = self._capture(n, primary[1], primary[2])   # state-neutral
            self._graphs[n] = primary               # _capture registered the new one under [n]
        entry = (self._graphs[n], self._graphs_alt[n])[self._parity]
        self._parity ^= 1
        graph, xs, ts, loss = entry
        cur = torch.cuda.current_stream(self.device)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        cs = self._copy_stream
        ready = self._db_events.get(id(graph))
        if ready is None:
            ready = self._db_events[id(graph)] = torch.cuda.Event()
        # Write-after-read on this input set: it was last read by the replay of TWO calls ago.
        # `_db_prev_launch` was recorded on the compute stream one call ago, right before that
        # call's replay was queued, i.e. after the replay of two calls ago: waiting for it makes the
        # copy safe and still lets it overlap the whole previous replay.  No such marker (first
        # call, or the synchronous path ran in between): wait for the compute stream as it is now.
        if isinstance(self._db_prev_launch, torch.cuda.Event):
            cs.wait_event(self._db_prev_launch)
        else:
            cs.wait_stream(cur)
        with torch.cuda.stream(cs):
            if xs is not None:
                xs.copy_(d, non_blocking=True)
            if ts is not None:
                ts.copy_(t, non_blocking=True)
            ready.record(cs)
        # two markers used alternately: the copy stream's wait above captured the PREVIOUS record of
        # the other one, so re-recording this one (last used two calls ago) cannot disturb it
        if self._db_markers is None:
            self._db_markers = [torch.cuda.Event(), torch.cuda.Event()]
        marker = self._db_markers[self._db_marker_idx]
        self._db_marker_idx ^= 1
        marker.record(cur)
        self._db_prev_launch = marker
        cur.wait_event(ready)
        graph.replay()
        return loss

    def train_batch(self, data: torch.Tensor, target: torch.Tensor, async_loss: bool = False):
        """PUBLIC per-step API: host tensors of one global batch in (pinned memory makes the
        H2D copy asynchronous), loss out.  Slices this replica's rows, copies them to the device,
        runs the step.  Returns the device loss tensor (``float(loss)`` is a synchronous D2H
        read), or with ``async_loss=True`` a :class:`LossHandle` whose D2H copy is already queued:
        read it after launching the next step to keep host and device overlapped (a handle stays
        valid for the next 7 steps: the pinned slots are a ring of 8)."""
        d, t, n = self._replica_slice(data, target)
        if n == 0:
            return None
        if self.graph_ok and n in self._graphs:
            if self._double_buffer and async_loss and d.is_pinned():
                loss = self._replay_double_buffered(n, d, t)
            else:
                graph, xs, ts, loss = self._graphs[n]
                if xs is not None:
                    xs.copy_(d, non_blocking=True)      # H2D straight into the graph's input
                if ts is not None:
                    ts.copy_(t, non_blocking=True)
                graph.replay()
                self._db_prev_launch = None
            self.stage.optimizer.step_count += 1
            self.engine.steps_done += 1
            if self.cfg.world_size > 1:
                if async_loss:
                    # the caller reads losses one step late, which already bounds the run-ahead;
                    # poll the pinned error words of the in-kernel deadlines now and then
                    self._peer_poll = (self._peer_poll + 1) & 15
                    if self._peer_poll == 0:
                        self.check_peers()
                else:
                    self._throttle()                    # bounded run-ahead + dead-peer check
            return self._async_loss(loss) if async_loss else loss
        x, tt = self._to_device(d, t)
        loss = self.train_step_device(x, tt, n // self.topo.tp)
        self._throttle()
        return self._async_loss(loss) if async_loss else loss

    # ---- loops (ref:106-117, ref:119-132) --------------------------------------------
    def train_epoch(self, epoch: int) -> None:
        cfg = self.cfg
        self.stage.set_training(True)
        n_batches = len(self.train_loader)
        t0 = time.perf_counter()
        steps_run = 0
        for batch_idx, (data, target) in enumerate(self.train_loader):
            if cfg.max_steps and batch_idx >= cfg.max_steps:
                break
            loss = self.train_batch(data, target)
            steps_run += 1
            self.global_step += 1
            self.samples_seen += data.shape[0]
            if batch_idx % cfg.log_interval == 0 and self.owns_loss:
                loss_v = float(loss.item())
                print(slog.train_line(epoch, batch_idx, data.shape[0], len(self.train_loader.dataset),
                                      n_batches, loss_v), file=self.out, flush=True)
                if self.metrics is not None:
                    self.metrics.write(kind="train", epoch=epoch, step=self.global_step, loss=loss_v,
                                       samples_seen=self.samples_seen)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
            self.check_peers()
        dt = time.perf_counter() - t0
        self.train_time_s += dt
        if self.metrics is not None:
            self.metrics.write(kind="epoch", epoch=epoch, seconds=dt, steps=steps_run,
                               samples_per_s=steps_run * cfg.batch_size * self.topo.dp / max(dt, 1e-9))

    def test_epoch(self) -> Optional[Tuple[float, int]]:
        if not getattr(self.stage, "supports_eval", True):
            return None
        self.stage.set_training(False)
        self.stage.eval_begin()
        for data, target in self.test_loader:
            # evaluation is not row-sharded over TP ranks (ragged last batch allowed)
            x, t = self._to_device(data, target, row_shard=False)
            self.engine.eval_batch(x, t, data.shape[0])
        if not self.stage.is_last:
            return None
        n = len(self.test_loader.dataset)
        test_loss = float(self.stage.loss_sum.item()) / n
        correct = int(self.stage.correct.item())
        if self.owns_loss:
            print(slog.test_block(test_loss, correct, n), file=self.out, flush=True)
            if self.metrics is not None:
                self.metrics.write(kind="test", loss=test_loss, correct=correct, total=n)
        return test_loss, correct

    def fit(self) -> None:
        cfg = self.cfg
        for epoch in range(self.start_epoch, cfg.epochs + 1):   # ref:134-136
            self.train_epoch(epoch)
            self.test_epoch()
            if cfg.ckpt_dir and (epoch == cfg.epochs or (cfg.save_every and epoch % cfg.save_every == 0)):
                self.save_checkpoint(epoch)

    # ---- checkpoint -------------------------------------------------------------------
    def save_checkpoint(self, epoch: int) -> None:
        # every replica collects (sharded DP state is gathered collectively), replica 0 writes
        ckpt.save_stage(self.cfg.ckpt_dir, self.stage, self.pp_idx, self.stage.optimizer,
                        self.tp_idx, self.topo.tp, write=self.dp_idx == 0)
        if self.rank == 0:
            ckpt.save_meta(self.cfg.ckpt_dir, epoch, self.global_step, self.cfg.asdict())
        self.barrier()

    def load_checkpoint(self) -> None:
        meta = ckpt.load_meta(self.cfg.ckpt_dir)
        if meta is None:
            return
        ckpt.load_stage(self.cfg.ckpt_dir, self.stage, self.pp_idx, self.stage.optimizer,
                        self.tp_idx, self.topo.tp)
        self.start_epoch = int(meta["epoch"]) + 1
        self.global_step = int(meta["step"])

    def barrier(self) -> None:
        if self.groups is not None:
            dist.barrier(group=self.groups.ctrl)

    def write_trace(self) -> Optional[dict]:
        """Dump this rank's timeline; rank 0 merges the files and prints the measured bubble."""
        if self.tracer is None:
            return None
        from .utils.trace import merge_traces
        self.tracer.dump(self.cfg.trace_dir)
        summary = self.tracer.summary()
        self.barrier()
        if self.rank == 0:
            merge_traces(self.cfg.trace_dir)
        if summary.get("steps"):
            print(slog.trace_line(summary), file=sys.stderr, flush=True)
        return summary

    def close_local(self) -> None:
        """Release what this Trainer holds on the device (native plans, p2p channels, peer mappings)
        but keep the process group: several Trainers can follow each other in one process."""
        for runner in self._runners.values():
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            runner.close()
        self._runners.clear()
        if getattr(self.engine, "native", None) is not None:
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            self.engine.native.close()
            self.engine.native = None
        if getattr(self.engine, "bnd", None) is not None:
            # peers may still be writing into this rank's slots: unmap only after everyone is done
            torch.cuda.synchronize(self.device)
            try:
                self.barrier()
            finally:
                self._graphs.clear()
                self._graphs_alt.clear()
                self.engine.bnd.close()
                self.engine.bnd = None
                if hasattr(self.stage, "boundary"):
                    self.stage.boundary = None

    def abort(self) -> None:
        """Tear-down after a failure (a peer died, a wait timed out): no barrier, no unmapping that a
        peer would have to take part in -- the process is about to end and the original exception must
        reach the user instead of a second time-out inside a barrier with dead ranks."""
        self._runners.clear()
        if self.metrics is not None:
            self.metrics.close()
        if dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:       # noqa: BLE001 -- best effort: the group may already be broken
                pass

    def close(self) -> None:
        try:
            self.write_trace()
        except Exception as e:      # a broken trace must not mask the training result
            print(f"[trace] rank {self.rank}: not written: {e}", file=sys.stderr)
        self.close_local()
        if self.metrics is not None:
            self.metrics.close()
        if dist.is_initialized():
            try:
                self.barrier()
            finally:
                dist.destroy_process_group()


def run(cfg: TrainConfig) -> Trainer:
    trainer = Trainer(cfg)
    try:
        trainer.fit()
    except BaseException:
        trainer.abort()   # do not wait in barriers for ranks that may be gone
        raise
    trainer.close()       # ref:186 rpc.shutdown(): graceful barrier, then tear down
    return trainer
