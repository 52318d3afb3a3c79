"""Wide-MLP stage: explicit forward / dgrad / wgrad equals autograd on the same math."""
import pytest
import torch

from simple_distributed_machine_learning_b200.models.mlp import MLPStage
from simple_distributed_machine_learning_b200.parallel import schedule as S
from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
from simple_distributed_machine_learning_b200.parallel.topology import Topology


def _autograd_reference(stages, x, targets, scale):
    """fp32 autograd over the same bf16-rounded weights/activations the stages use."""
    ws, bs = [], []
    for st in stages:
        for l in range(st.n_layers):
            ws.append(st.w[l].float().clone().requires_grad_(True))
            bs.append(st.b_master[l].clone().requires_grad_(True))
    h = x.to(torch.bfloat16).float()
    n = len(ws)
    for i, (w, b) in enumerate(zip(ws, bs)):
        h = h @ w.t() + b
        if i < n - 1:
            h = torch.relu(h)
        h = h.detach().to(torch.bfloat16).float() + (h - h.detach())  # straight-through bf16 rounding
    loss = torch.nn.functional.cross_entropy(h, targets, reduction="sum") * scale
    loss.backward()
    return loss.item(), [w.grad for w in ws], [b.grad for b in bs]


def _run_local_pipeline(stages, x, t, m, kind):
    """Drive p stages in one process following the generated op lists (no comm)."""
    p = len(stages)
    B = x.shape[0]
    sizes = [B // m] * m
    offs = [i * (B // m) for i in range(m + 1)]
    ops = S.generate_all(kind, p, m)
    pc = [0] * p
    acts, grads = {}, {}
    for st in stages:
        st.begin_step(B, sizes)
    progressed = True
    while progressed:
        progressed = False
        for s, st in enumerate(stages):
            while pc[s] < len(ops[s]):
                op = ops[s][pc[s]]
                if op.kind == S.OpKind.F:
                    if s > 0 and (s - 1, op.mb) not in acts:
                        break
                    xin = x[offs[op.mb]:offs[op.mb + 1]] if s == 0 else acts.pop((s - 1, op.mb))
                    y = st.forward(op.mb, xin, t[offs[op.mb]:offs[op.mb + 1]] if st.is_last else None)
                    if not st.is_last:
                        acts[(s, op.mb)] = y
                elif op.kind in (S.OpKind.B, S.OpKind.BW):
                    if s < p - 1 and (s + 1, op.mb) not in grads:
                        break
                    dy = None if st.is_last else grads.pop((s + 1, op.mb))
                    dx = st.backward_input(op.mb, dy)
                    if not st.is_first:
                        grads[(s, op.mb)] = dx
                    if op.kind == S.OpKind.BW:
                        st.backward_weight(op.mb)
                else:
                    st.backward_weight(op.mb)
                pc[s] += 1
                progressed = True
    assert all(pc[s] == len(ops[s]) for s in range(p))
    for st in stages:
        st.flush_weight_grads()


@pytest.mark.parametrize("kind,p,m", [("1f1b", 2, 4), ("zb-h1", 2, 4), ("zb-h2", 4, 4), ("gpipe", 1, 2)])
def test_manual_backward_matches_autograd(kind, p, m):
    torch.manual_seed(0)
    H, C, B, L = 64, 32, 32, 4
    dev = torch.device("cpu")
    stages = [MLPStage(s, p, L, H, C, dev, lr=0.0, momentum=0.0) for s in range(p)]
    x = torch.randn(B, H)
    t = torch.randint(0, C, (B,))
    ref_loss, ref_dw, ref_db = _autograd_reference(stages, x, t, 1.0 / B)
    _run_local_pipeline(stages, x, t, m, kind)
    got_dw = [g for st in stages for g in st.dw]
    got_db = [g for st in stages for g in st.db]
    loss = stages[-1].mean_loss().item()
    assert loss == pytest.approx(ref_loss, rel=2e-2)
    for a, b in zip(got_dw, ref_dw):
        torch.testing.assert_close(a, b, rtol=5e-2, atol=2e-3)
    for a, b in zip(got_db, ref_db):
        torch.testing.assert_close(a, b, rtol=5e-2, atol=2e-3)


def test_mlp_training_reduces_loss_and_checkpoint_names():
    torch.manual_seed(0)
    H, C, B = 64, 16, 64
    dev = torch.device("cpu")
    st = MLPStage(0, 1, 2, H, C, dev, lr=0.05, momentum=0.5)
    eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=4, schedule="zb-h1")
    x = torch.randn(B, H)
    t = torch.randint(0, C, (B,))
    losses = [eng.train_step(x, t, B).item() for _ in range(30)]
    assert losses[-1] < 0.5 * losses[0]
    sd = st.state_dict()
    assert sorted(sd) == ["layers.0.bias", "layers.0.weight", "layers.1.bias", "layers.1.weight"]
    seq = torch.nn.Sequential()  # an nn.Module with the same names can load it
    holder = torch.nn.Module()
    holder.layers = torch.nn.ModuleList([torch.nn.Linear(H, H), torch.nn.Linear(H, C)])
    holder.load_state_dict(sd, strict=True)
    # bf16 compute copies follow the fp32 masters
    assert torch.equal(st.w[0], st.w_master[0].to(torch.bfloat16))


def test_fp8_forward_and_backward_plumbing_on_cpu(monkeypatch):
    """dtype fp8 with the CPU emulation of the MX-fp8 kernels (same arithmetic, same byte layouts):
    dual row-wise / transposed copies, step-wide transposed buffers filled per micro-batch, K-slices
    of them as wgrad operands, fp32 accumulation across wgrad groups -- across two pipeline stages.

    The quantised arithmetic does not depend on the schedule or on how micro-batches are grouped
    into wgrad GEMMs, so every (schedule, group size) must give the same gradients up to fp32
    summation order; against the bf16 stages only a loose bound holds (a few per cent of the ReLU
    masks flip when activations carry fp8 noise, which dominates the gradient difference)."""
    monkeypatch.setenv("SDML_FP8_BWD", "1")
    H, C, B, L, m = 256, 128, 512, 4, 4
    x = torch.randn(B, H, generator=torch.Generator().manual_seed(0))
    t = torch.randint(0, C, (B,), generator=torch.Generator().manual_seed(1))

    def run(dtype, kind, w_group):
        stages = [MLPStage(s, 2, L, H, C, torch.device("cpu"), lr=0.0, momentum=0.0, dtype=dtype) for s in range(2)]
        for st in stages:
            st.w_group = w_group
        _run_local_pipeline(stages, x, t, m, kind)
        if dtype == "fp8":
            assert all(st.fp8_backward and all(st._fp8_bwd_ok) for st in stages)
            from simple_distributed_machine_learning_b200.ops import fp8 as fp8_mod
            used = all(len(st._xq_ready) == m * (st.n_layers - 1) and len(st._dzq_ready) == m * (st.n_layers - 1)
                       for st in stages)
            assert used == bool(fp8_mod._FUSED_QOUT), "producer-side quantisation used / not used as configured"
        return ([g.clone() for st in stages for g in st.dw], [g.clone() for st in stages for g in st.db],
                stages[-1].mean_loss().item())

    def run_plain(kind, w_group):
        from simple_distributed_machine_learning_b200.ops import fp8 as fp8_mod
        old = fp8_mod._FUSED_QOUT
        fp8_mod._FUSED_QOUT = False
        try:
            return run("fp8", kind, w_group)
        finally:
            fp8_mod._FUSED_QOUT = old

    base = run("fp8", "gpipe", 0)               # one wgrad GEMM per layer over the whole step
    for kind, w_group in (("zb-h1", 1), ("1f1b", 2), ("zb-h2", 3)):
        other = run("fp8", kind, w_group)
        assert other[2] == pytest.approx(base[2], rel=1e-6)
        for a, b in zip(other[0] + other[1], base[0] + base[1]):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-7)
    # producer-side quantisation (the GEMM that writes an activation / gradient also emits its fp8
    # copies; on GPUs from its epilogue): bit-identical to the consumer-side quantiser passes
    from simple_distributed_machine_learning_b200.ops import fp8 as fp8_mod
    monkeypatch.setattr(fp8_mod, "_FUSED_QOUT", True)
    for kind, w_group in (("gpipe", 0), ("zb-h1", 1)):
        fused = run("fp8", kind, w_group)
        ref = base if w_group == 0 else run_plain(kind, w_group)
        assert fused[2] == ref[2]
        for a, b in zip(fused[0] + fused[1], ref[0] + ref[1]):
            assert torch.equal(a, b)
    monkeypatch.setattr(fp8_mod, "_FUSED_QOUT", False)
    bf16 = run("bf16", "gpipe", 0)
    assert base[2] == pytest.approx(bf16[2], rel=2e-2)
    errs = [((g8 - g16).norm() / g16.norm()).item() for g8, g16 in zip(base[0], bf16[0])]
    assert errs[-1] < 0.08 and max(errs) < 0.5, errs      # last layer: pure quantisation error


def test_fp8_training_on_cpu_tracks_bf16(monkeypatch):
    monkeypatch.setenv("SDML_FP8_BWD", "1")
    g = torch.Generator().manual_seed(5)
    x = torch.randn(512, 256, generator=g)
    t = torch.randint(0, 128, (512,), generator=g)
    curves = {}
    for dtype in ("bf16", "fp8"):
        st = MLPStage(0, 1, 3, 256, 128, torch.device("cpu"), lr=0.02, momentum=0.5, dtype=dtype)
        eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=2, schedule="gpipe")
        curves[dtype] = [eng.train_step(x, t, 512).item() for _ in range(15)]
    assert curves["fp8"][-1] < 0.9 * curves["fp8"][0]
    for a, b in zip(curves["fp8"], curves["bf16"]):
        assert a == pytest.approx(b, rel=0.06), (curves["fp8"][::3], curves["bf16"][::3])


def test_fp8_copies_follow_loaded_weights(monkeypatch):
    """load_state_dict after the stage has already run: the quantised weight copies are refreshed."""
    monkeypatch.setenv("SDML_FP8_BWD", "1")
    x = torch.randn(256, 256, generator=torch.Generator().manual_seed(0))
    t = torch.randint(0, 128, (256,), generator=torch.Generator().manual_seed(1))
    a = MLPStage(0, 1, 2, 256, 128, torch.device("cpu"), lr=0.05, momentum=0.0, dtype="fp8", seed=1)
    b = MLPStage(0, 1, 2, 256, 128, torch.device("cpu"), lr=0.05, momentum=0.0, dtype="fp8", seed=2)
    ea = PipelineEngine(a, Topology(1, 1), 0, None, microbatches=2, schedule="gpipe")
    eb = PipelineEngine(b, Topology(1, 1), 0, None, microbatches=2, schedule="gpipe")
    ea.train_step(x, t, 256)
    eb.train_step(x, t, 256)                 # b's fp8 copies now exist and belong to ITS weights
    b.load_state_dict(a.state_dict())
    la, lb = ea.train_step(x, t, 256).item(), eb.train_step(x, t, 256).item()
    assert la == lb


def test_buffers_of_a_batch_size_survive_another_batch_size():
    """A CUDA graph captured for the full batch size replays into the stage's row-stacked buffers; a
    ragged last batch of an epoch must switch to its own buffer set, not free the first one."""
    import torch
    from simple_distributed_machine_learning_b200.models.mlp import MLPStage
    st = MLPStage(0, 1, 2, 32, 16, torch.device("cpu"), lr=0.01)
    st.begin_step(64, [32, 32])
    first = [t.data_ptr() for t in st._X + st._dZ + [st._Y]]
    keep = st._X[0]
    st.begin_step(40, [20, 20])
    assert st._X[0].shape[0] == 40 and st._X[0].data_ptr() != keep.data_ptr()
    second = [t.data_ptr() for t in st._X + st._dZ + [st._Y]]
    st.begin_step(64, [32, 32])
    assert [t.data_ptr() for t in st._X + st._dZ + [st._Y]] == first
    st.begin_step(40, [20, 20])
    assert [t.data_ptr() for t in st._X + st._dZ + [st._Y]] == second
