"""Tensor-parallel MLP stage over gloo on CPUs: the column-/row-parallel layer pairs with
all-gather / reduce-scatter (the NCCL-style baseline backend of TPMLPStage) must train to the
same weights as the unsharded stage.  The fused GPU kernels are checked against the same
baseline in tests/test_tp_gpu.py and tests/test_multigpu.py."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from tests import _workers as W

CFG = dict(model="mlp", hidden=256, layers=2, classes=256, dtype="bf16", batch_size=512, epochs=1,
           train_samples=1024, test_samples=64, lr=0.02, momentum=0.5, log_interval=1, microbatches=2,
           schedule="gpipe")


def _spawn(world, port, kw, out_dir, timeout=240):
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=W.train_worker, args=(r, world, port, kw, out_dir)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    assert all(not p.is_alive() for p in procs), "worker hung"
    assert [p.exitcode for p in procs] == [0] * world


def _load(d, world):
    return [torch.load(os.path.join(d, f"rank{r}.pt"), weights_only=False) for r in range(world)]


@pytest.fixture(scope="module")
def single(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("tp_single"))
    W.train_worker(0, 1, 0, dict(CFG), d)
    return _load(d, 1)[0]


def _test_line(log):
    line = [l for l in log.splitlines() if l.startswith("Test set:")][-1]
    loss = float(line.split("Average loss: ")[1].split(",")[0])
    correct = int(line.split("Accuracy: ")[1].split("/")[0])
    return loss, correct


def _full_from_shards(shards, pair=0):
    a, b = 2 * pair, 2 * pair + 1
    wa = torch.cat([s[f"layers.{a}.weight"] for s in shards], 0)
    ba = torch.cat([s[f"layers.{a}.bias"] for s in shards], 0)
    wb = torch.cat([s[f"layers.{b}.weight"] for s in shards], 1)
    return wa, ba, wb, shards[0][f"layers.{b}.bias"]


def test_tp2_matches_unsharded(tmp_path, free_port, single):
    _spawn(2, free_port, dict(CFG, pp=1, tp=2), str(tmp_path))
    res = _load(str(tmp_path), 2)
    shards = [r["parts"][0] for r in sorted(res, key=lambda r: r["coords"][2])]
    wa, ba, wb, bb = _full_from_shards(shards)
    # the evaluation pass (ref:119-132) runs on TP stages too: same test loss as the unsharded model
    owner = [r for r in res if r["coords"][2] == 0][0]
    loss_tp, _ = _test_line(owner["log"])
    loss_1, _ = _test_line(single["log"])
    assert loss_tp == pytest.approx(loss_1, rel=2e-2)
    single = single["parts"][0]
    torch.testing.assert_close(wa, single["layers.0.weight"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(ba, single["layers.0.bias"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(wb, single["layers.1.weight"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(bb, single["layers.1.bias"], rtol=2e-2, atol=2e-3)


def test_pp2_tp2_matches_unsharded(tmp_path, free_port, tmp_path_factory):
    cfg = dict(CFG, layers=4)
    d1 = str(tmp_path_factory.mktemp("tp_single4"))
    W.train_worker(0, 1, 0, dict(cfg), d1)
    ref = _load(d1, 1)[0]["parts"][0]
    _spawn(4, free_port, dict(cfg, pp=2, tp=2, schedule="1f1b"), str(tmp_path))
    res = _load(str(tmp_path), 4)
    for stage in (0, 1):
        shards = [r["parts"][stage] for r in sorted(res, key=lambda r: r["coords"][2]) if r["coords"][0] == stage]
        wa, ba, wb, bb = _full_from_shards(shards)
        base = 2 * stage
        torch.testing.assert_close(wa, ref[f"layers.{base}.weight"], rtol=3e-2, atol=3e-3)
        torch.testing.assert_close(wb, ref[f"layers.{base + 1}.weight"], rtol=3e-2, atol=3e-3)
        torch.testing.assert_close(bb, ref[f"layers.{base + 1}.bias"], rtol=3e-2, atol=3e-3)


def test_tp2_fp8_matches_unsharded_fp8(tmp_path, free_port, tmp_path_factory):
    """dtype fp8 in the tensor-parallel stage (BASELINE config 5's dtype): every GEMM of the stage on
    the block-scaled path (CPU emulation here), row-wise + transposed copies of the gathered
    operands.  MX blocks never straddle a shard (hidden/tp is a multiple of 128), so the sharded
    arithmetic equals the unsharded fp8 stage's up to summation order."""
    cfg = dict(CFG, dtype="fp8")
    d1 = str(tmp_path_factory.mktemp("tp_fp8_single"))
    W.train_worker(0, 1, 0, dict(cfg), d1)
    ref = _load(d1, 1)[0]
    _spawn(2, free_port, dict(cfg, pp=1, tp=2), str(tmp_path))
    res = _load(str(tmp_path), 2)
    shards = [r["parts"][0] for r in sorted(res, key=lambda r: r["coords"][2])]
    wa, ba, wb, bb = _full_from_shards(shards)
    single = ref["parts"][0]
    torch.testing.assert_close(wa, single["layers.0.weight"], rtol=3e-2, atol=3e-3)
    torch.testing.assert_close(wb, single["layers.1.weight"], rtol=3e-2, atol=3e-3)
    torch.testing.assert_close(bb, single["layers.1.bias"], rtol=3e-2, atol=3e-3)
    owner = [r for r in res if r["coords"][2] == 0][0]
    assert _test_line(owner["log"])[0] == pytest.approx(_test_line(ref["log"])[0], rel=3e-2)
    # and the fp8 run is close to the bf16 one (same data, same init)
    d2 = str(tmp_path_factory.mktemp("tp_bf16_single"))
    W.train_worker(0, 1, 0, dict(CFG), d2)
    assert _test_line(owner["log"])[0] == pytest.approx(_test_line(_load(d2, 1)[0]["log"])[0], rel=5e-2)


def test_pp2_tp2_fp8_trains_like_bf16(tmp_path, free_ports, tmp_path_factory):
    """BASELINE config 5's shape in miniature: pipeline stages x 2-way tensor parallel with block-scaled
    fp8 GEMMs (CPU emulation of the kernels): the test loss after one epoch tracks the bf16 run."""
    cfg = dict(CFG, layers=4, pp=2, tp=2, schedule="zb-h1")
    out = {}
    ports = free_ports(2)
    for i, dtype in enumerate(("bf16", "fp8")):
        d = str(tmp_path_factory.mktemp(f"pptp_{dtype}"))
        _spawn(4, ports[i], dict(cfg, dtype=dtype), d)
        res = _load(d, 4)
        owner = [r for r in res if r["coords"][0] == 1 and r["coords"][2] == 0][0]
        out[dtype] = _test_line(owner["log"])[0]
    assert out["fp8"] == pytest.approx(out["bf16"], rel=5e-2)


def test_tp_shard_checkpoints_resume_exactly(tmp_path, free_ports):
    """One file per pipeline stage AND tensor-parallel shard (`stageN_tpK.pt` + optimizer state);
    two epochs in one go == one epoch, checkpoint, resume for the second."""
    cfg = dict(CFG, pp=1, tp=2, epochs=2)
    a, b, ck = tmp_path / "a", tmp_path / "b", tmp_path / "ck"
    for d in (a, b):
        d.mkdir()
    p0, p1, p2 = free_ports(3)
    _spawn(2, p0, dict(cfg), str(a))                                               # 2 epochs straight
    _spawn(2, p1, dict(cfg, epochs=1, ckpt_dir=str(ck)), str(b))                   # epoch 1, save
    assert sorted(os.listdir(ck)) == ["meta.json", "stage0_tp0.optim.pt", "stage0_tp0.pt",
                                      "stage0_tp1.optim.pt", "stage0_tp1.pt"]
    _spawn(2, p2, dict(cfg, epochs=2, ckpt_dir=str(ck), resume=True), str(b))      # resume -> epoch 2
    ra, rb = _load(str(a), 2), _load(str(b), 2)
    for x, y in zip(sorted(ra, key=lambda r: r["coords"][2]), sorted(rb, key=lambda r: r["coords"][2])):
        for k in x["parts"][0]:
            torch.testing.assert_close(x["parts"][0][k], y["parts"][0][k], rtol=0, atol=0)
