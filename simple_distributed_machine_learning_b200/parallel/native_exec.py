"""Python face of the native op-list executor (``csrc/executor.cu``).

Two things live here:

* :func:`run_with_hooks` -- the C++ control loop driven by Python callbacks (used by the CPU
  tests to check the order of actions, and usable with any backend);
* :class:`NativeStepRunner` -- the GPU binding (``SDML_EXECUTOR=native``): every op of this
  rank's schedule is captured ONCE into its own CUDA graph, stage-boundary transfers go through
  the native NCCL engine, and a training step is ONE call into C++ (`sdml_exec_plan_step`)
  that issues every graph launch, receive, wait and send.  The reference spends four blocking
  Python statements per batch (/root/reference/simple_distributed.py:108-113).

Requirements of the GPU binding (checked, else the Python executor keeps running): CUDA, SGD,
no DP / TP inside the stage (their reductions would have to be captured too), a stage with
static per-micro-batch receive slots (``input_buffer`` / ``grad_buffer``), and -- for more than
one stage -- the native p2p channels (``SDML_P2P=native``).
"""
from __future__ import annotations

import ctypes
from typing import Callable, List, Optional, Sequence

import torch

from ..ops import _lib
from .schedule import Op, OpKind

_HOOK_COMPUTE = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int)
_HOOK_POST = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                              ctypes.POINTER(ctypes.c_longlong))
_HOOK_WAIT = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong)
_HOOK_END = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p)


class _Hooks(ctypes.Structure):
    _fields_ = [("user", ctypes.c_void_p), ("compute", _HOOK_COMPUTE), ("post_recv", _HOOK_POST),
                ("wait_recv", _HOOK_WAIT), ("send", _HOOK_POST), ("wait_send", _HOOK_WAIT),
                ("end_step", _HOOK_END)]


def _declare(lib) -> None:
    if getattr(lib, "_sdml_exec_declared", False):
        return
    I, P, L = ctypes.c_int, ctypes.c_void_p, ctypes.c_ulonglong
    sigs = {
        "sdml_exec_run_hooks": [ctypes.POINTER(I), I, I, I, I, ctypes.POINTER(_Hooks)],
        "sdml_exec_plan_create": [ctypes.POINTER(I), I, I, I, I, P, ctypes.POINTER(P)],
        "sdml_exec_plan_set_graph": [P, I, I, P],
        "sdml_exec_plan_set_channel": [P, I, I, P],
        "sdml_exec_plan_set_buffer": [P, I, I, I, P, L],
        "sdml_exec_plan_step": [P],
        "sdml_exec_plan_destroy": [P],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.restype = I
        fn.argtypes = argtypes
    lib._sdml_exec_declared = True


def _ops_array(ops: Sequence[Op]):
    arr = (ctypes.c_int * (2 * len(ops)))()
    for i, op in enumerate(ops):
        arr[2 * i], arr[2 * i + 1] = int(op.kind), op.mb
    return arr


def run_with_hooks(ops: Sequence[Op], is_first: bool, is_last: bool, lookahead: int, *,
                   compute: Callable[[int, int], None], post_recv: Callable[[int, int], int],
                   wait_recv: Callable[[int, int, int], None], send: Callable[[int, int], int],
                   wait_send: Callable[[int, int, int], None], end_step: Callable[[], None]) -> None:
    """Run one step of ``ops`` through the C++ loop; the hooks are plain Python callables
    (``post_recv`` / ``send`` return a ticket).  Exceptions raised by a hook abort the step."""
    lib = _lib.cuda_lib()
    _declare(lib)
    err: List[BaseException] = []

    def guard(fn):
        def wrapped(*a):
            try:
                r = fn(*a)
                return 0 if r is None else r
            except BaseException as e:  # noqa: BLE001 -- must not propagate through C frames
                err.append(e)
                return -2
        return wrapped

    def post(fn):
        def wrapped(_user, d, mb, out):
            out[0] = int(fn(d, mb))
            return 0
        return guard(wrapped)

    hooks = _Hooks(None,
                   _HOOK_COMPUTE(guard(lambda _u, k, mb: compute(k, mb))),
                   _HOOK_POST(post(post_recv)),
                   _HOOK_WAIT(guard(lambda _u, d, mb, t: wait_recv(d, mb, t))),
                   _HOOK_POST(post(send)),
                   _HOOK_WAIT(guard(lambda _u, d, mb, t: wait_send(d, mb, t))),
                   _HOOK_END(guard(lambda _u: end_step())))
    rc = lib.sdml_exec_run_hooks(_ops_array(ops), len(ops), int(is_first), int(is_last), lookahead,
                                 ctypes.byref(hooks))
    if err:
        raise err[0]
    _lib.check(rc, "sdml_exec_run_hooks")


class NativeStepRunner:
    """Per-op CUDA graphs + native p2p, one C++ call per training step."""

    @staticmethod
    def supported(engine) -> Optional[str]:
        """None when the engine can be driven natively, else the reason why not."""
        st = engine.stage
        if not engine.is_cuda:
            return "not on CUDA"
        if engine.topo.dp != 1 or engine.topo.tp != 1:
            return "DP / TP reductions are not captured"
        if type(st.optimizer).__name__ != "FusedSGD":
            return "only the fused SGD step is capturable (Adam's bias correction is a host scalar)"
        if engine.p > 1 and engine.native is None:
            return "needs the native p2p channels (SDML_P2P=native)"
        if engine.tracer is not None:
            return "tracing runs the step eagerly"
        return None

    def __init__(self, engine, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor],
                 batch_size: int):
        from ..utils.data import split_microbatches
        self.engine = engine
        st = self.stage = engine.stage
        dev = st.device
        lib = self.lib = _lib.cuda_lib()
        _declare(lib)
        sizes = split_microbatches(batch_size, engine.m)
        n_mb = len(sizes)
        offs = [0]
        for s in sizes:
            offs.append(offs[-1] + s)
        ops = engine.ops_for(n_mb)
        self.batch_size, self.sizes = batch_size, sizes
        # static step inputs (the caller's batch is copied in before every step)
        self.xs = torch.empty_like(inputs) if (st.is_first and inputs is not None) else None
        self.ts = torch.empty_like(targets) if (st.is_last and targets is not None) else None
        stream = torch.cuda.current_stream(dev)
        plan = ctypes.c_void_p()
        _lib.check(lib.sdml_exec_plan_create(_ops_array(ops), len(ops), n_mb, int(st.is_first), int(st.is_last),
                                             stream.cuda_stream or None, ctypes.byref(plan)), "sdml_exec_plan_create")
        self._plan = plan
        self._graphs: List[torch.cuda.CUDAGraph] = []
        self._keep: List[torch.Tensor] = []       # everything captured ops handed out stays alive
        pool = torch.cuda.graph_pool_handle()

        def capture(kind: int, mb: int, fn):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=pool, stream=self._cap_stream):
                out = fn()
                st.join_side_streams()      # a capture must end with every forked stream joined
            self._graphs.append(g)
            _lib.check(lib.sdml_exec_plan_set_graph(plan, kind, mb, g.raw_cuda_graph_exec()), "set_graph")
            return out

        def set_buf(direction: int, is_send: bool, mb: int, t: torch.Tensor):
            assert t.is_contiguous()
            self._keep.append(t)
            _lib.check(lib.sdml_exec_plan_set_buffer(plan, direction, int(is_send), mb, t.data_ptr(),
                                                     t.numel() * t.element_size()), "set_buffer")

        # Walk the op list once in capture mode: host-side bookkeeping of the stage (pending wgrad
        # groups, accumulate flags) evolves exactly as in a real step, so every op's graph holds the
        # kernels that op launches AT THAT POSITION of the schedule.  Nothing executes.
        self._cap_stream = torch.cuda.Stream(device=dev)
        self._cap_stream.wait_stream(stream)
        opt = st.optimizer
        saved_counts = (opt.step_count, engine.steps_done)
        self._begin = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._begin, pool=pool, stream=self._cap_stream):
            st.begin_step(batch_size, sizes)
        engine._in_step = True
        for op in ops:
            mb = op.mb
            if op.kind == OpKind.F:
                if st.is_first:
                    x = self.xs[offs[mb]:offs[mb + 1]]
                else:
                    x = st.input_buffer(mb)
                    if x is None:
                        raise RuntimeError("stage has no static receive slot for activations")
                    set_buf(0, False, mb, x)
                t = self.ts[offs[mb]:offs[mb + 1]] if st.is_last else None
                y = capture(0, mb, lambda: st.forward(mb, x, t))
                if not st.is_last:
                    set_buf(0, True, mb, y)
            elif op.kind in (OpKind.B, OpKind.BW):
                dy = None
                if not st.is_last:
                    dy = st.grad_buffer(mb)
                    if dy is None:
                        raise RuntimeError("stage has no static receive slot for gradients")
                    set_buf(1, False, mb, dy)
                dx = capture(1, mb, lambda: st.backward_input(mb, dy))
                if not st.is_first:
                    set_buf(1, True, mb, dx)
                if op.kind == OpKind.BW:
                    capture(2, mb, lambda: st.backward_weight(mb))
            else:
                capture(2, mb, lambda: st.backward_weight(mb))

        def end():
            st.end_step()
            return st.mean_loss() if st.is_last else None
        self.loss = capture(3, 0, end)
        engine._in_step = False
        opt.step_count, engine.steps_done = saved_counts     # the captured walk was not a step
        stream.wait_stream(self._cap_stream)
        if engine.native is not None:
            ch = engine.native
            for direction, is_send, c in ((0, False, ch.act_in), (1, False, ch.grad_in),
                                          (0, True, ch.act_out), (1, True, ch.grad_out)):
                if c is not None:
                    _lib.check(lib.sdml_exec_plan_set_channel(plan, direction, int(is_send), c._h), "set_channel")

    def step(self, inputs: Optional[torch.Tensor], targets: Optional[torch.Tensor]):
        """One optimizer step; returns the mean loss (static device tensor) on the last stage."""
        if self.xs is not None:
            self.xs.copy_(inputs, non_blocking=True)
        if self.ts is not None:
            self.ts.copy_(targets, non_blocking=True)
        self._begin.replay()
        _lib.check(self.lib.sdml_exec_plan_step(self._plan), "sdml_exec_plan_step")
        self.stage.optimizer.step_count += 1
        self.engine.steps_done += 1
        return self.loss

    def close(self) -> None:
        if self._plan:
            self.lib.sdml_exec_plan_destroy(self._plan)
            self._plan = ctypes.c_void_p()
