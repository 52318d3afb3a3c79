"""The C++ op-list executor (csrc/executor.cu).

CPU part: the control loop, driven through Python hooks, must issue exactly the actions of the
Python executor (parallel/pipeline.py) in the same order for every schedule / rank / look-ahead.
GPU part: per-op CUDA graphs + one C++ call per step reproduce the Python executor's training."""
import pytest
import torch

from simple_distributed_machine_learning_b200.parallel import schedule as sched
from simple_distributed_machine_learning_b200.parallel.schedule import OpKind


def _python_reference(ops, is_first, is_last, lookahead):
    """The loop of PipelineEngine.train_step, reduced to its action trace."""
    log, ticket = [], [0]

    def new_ticket():
        ticket[0] += 1
        return ticket[0]
    needs = [i for i, o in enumerate(ops)
             if (o.kind == OpKind.F and not is_first) or (o.kind in (OpKind.B, OpKind.BW) and not is_last)]
    posted, sends, nxt = {}, [], 0
    for i, op in enumerate(ops):
        while nxt < len(needs) and needs[nxt] <= i + lookahead:
            j = needs[nxt]
            d = 0 if ops[j].kind == OpKind.F else 1
            posted[j] = (d, ops[j].mb, new_ticket())
            log.append(("post_recv", d, ops[j].mb))
            nxt += 1
        if i in posted:
            d, mb, t = posted.pop(i)
            log.append(("wait_recv", d, mb, t))
        if op.kind == OpKind.F:
            log.append(("compute", 0, op.mb))
            if not is_last:
                sends.append((0, op.mb, new_ticket()))
                log.append(("send", 0, op.mb))
        elif op.kind in (OpKind.B, OpKind.BW):
            log.append(("compute", 1, op.mb))
            if not is_first:
                sends.append((1, op.mb, new_ticket()))
                log.append(("send", 1, op.mb))
            if op.kind == OpKind.BW:
                log.append(("compute", 2, op.mb))
        else:
            log.append(("compute", 2, op.mb))
    log.append(("end_step",))
    log.extend(("wait_send", d, mb, t) for d, mb, t in sends)
    return log


@pytest.mark.parametrize("kind", ["gpipe", "1f1b", "zb-h1", "zb-h2"])
@pytest.mark.parametrize("p,m", [(1, 4), (2, 8), (4, 8), (8, 16)])
@pytest.mark.parametrize("lookahead", [0, 2])
def test_cpp_loop_issues_the_python_executors_actions(kind, p, m, lookahead):
    from simple_distributed_machine_learning_b200.parallel.native_exec import run_with_hooks
    for rank in range(p):
        ops = sched.generate(kind, p, m, rank)
        log, ticket = [], [0]

        def new_ticket():
            ticket[0] += 1
            return ticket[0]
        run_with_hooks(
            ops, rank == 0, rank == p - 1, lookahead,
            compute=lambda k, mb: log.append(("compute", k, mb)),
            post_recv=lambda d, mb: (log.append(("post_recv", d, mb)), new_ticket())[1],
            wait_recv=lambda d, mb, t: log.append(("wait_recv", d, mb, t)),
            send=lambda d, mb: (log.append(("send", d, mb)), new_ticket())[1],
            wait_send=lambda d, mb, t: log.append(("wait_send", d, mb, t)),
            end_step=lambda: log.append(("end_step",)))
        assert log == _python_reference(ops, rank == 0, rank == p - 1, lookahead), (kind, p, m, rank)


def test_hook_exception_aborts_the_step_and_propagates():
    from simple_distributed_machine_learning_b200.parallel.native_exec import run_with_hooks
    ops = sched.generate("1f1b", 2, 4, 0)
    seen = []

    def compute(k, mb):
        seen.append((k, mb))
        if len(seen) == 3:
            raise ValueError("boom")
    with pytest.raises(ValueError, match="boom"):
        run_with_hooks(ops, True, False, 0, compute=compute, post_recv=lambda d, mb: 1,
                       wait_recv=lambda d, mb, t: None, send=lambda d, mb: 1,
                       wait_send=lambda d, mb, t: None, end_step=lambda: None)
    assert len(seen) == 3          # nothing ran after the failing hook


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,schedule", [("bf16", "zb-h1"), ("fp8", "gpipe")])
def test_native_runner_matches_python_executor_on_one_gpu(dtype, schedule):
    """pp = 1: no transfers, but every F / B / W op and the optimizer step run as separately
    captured CUDA graphs launched from C++; losses must match the Python executor's."""
    from simple_distributed_machine_learning_b200.models.mlp import MLPStage
    from simple_distributed_machine_learning_b200.parallel.native_exec import NativeStepRunner
    from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
    from simple_distributed_machine_learning_b200.parallel.topology import Topology
    g = torch.Generator().manual_seed(7)
    x = torch.randn(1024, 512, generator=g).cuda().bfloat16()
    t = torch.randint(0, 256, (1024,), generator=g).cuda()
    curves = {}
    for mode in ("python", "native"):
        st = MLPStage(0, 1, 3, 512, 256, torch.device("cuda"), lr=0.02, momentum=0.5, dtype=dtype)
        st.w_group = 2
        eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=4, schedule=schedule)
        losses = [eng.train_step(x, t, 1024).item() for _ in range(3)]
        if mode == "native":
            assert NativeStepRunner.supported(eng) is None
            runner = NativeStepRunner(eng, x, t, 1024)
            losses += [runner.step(x, t).item() for _ in range(9)]
            assert eng.steps_done == 12 and st.optimizer.step_count == 12
            runner.close()
        else:
            losses += [eng.train_step(x, t, 1024).item() for _ in range(9)]
        curves[mode] = losses
    assert curves["native"][-1] < 0.9 * curves["native"][0]
    for a, b in zip(curves["native"], curves["python"]):
        assert a == pytest.approx(b, rel=2e-3), (curves["native"], curves["python"])
