"""Multi-GPU paths (NCCL p2p pipeline, fused in-kernel DP all-reduce + SGD): each layout
must reproduce the single-GPU run from identical initial weights.  Needs >= 2 GPUs
(`gpurun --gpus 2 -- python -m pytest tests/test_multigpu.py -m gpu`)."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from tests import _workers as W

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _spawn(fn, world, port, kw, out_dir, timeout=180):
    timeout = float(os.environ.get("SDML_TEST_SPAWN_TIMEOUT", timeout))
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=fn, args=(r, world, port, kw, out_dir)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.kill()
    assert not hung, "worker hung"
    assert [p.exitcode for p in procs] == [0] * world


def _load(d, world):
    return [torch.load(os.path.join(d, f"rank{r}.pt"), weights_only=False) for r in range(world)]


def _merged(results):
    parts = {}
    for r in results:
        if r["coords"][1] == 0 and r["coords"][2] == 0:
            parts.update(r["parts"])
    return parts


def _close(a, b, tol):
    assert sorted(a) == sorted(b)
    for idx in a:
        for k in a[idx]:
            torch.testing.assert_close(a[idx][k], b[idx][k], rtol=tol, atol=tol, msg=lambda m: f"stage{idx}.{k}: {m}")


LENET = dict(epochs=1, train_samples=480, test_samples=120, dropout=False, log_interval=4)


@pytest.fixture(scope="module")
def lenet_single(tmp_path_factory):
    if _ngpu() < 1:
        pytest.skip("no GPU")
    d = str(tmp_path_factory.mktemp("single"))
    _spawn(W.gpu_train_worker, 1, 29871, dict(LENET), d)
    return _load(d, 1)[0]


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("schedule,mb", [("1f1b", 8), ("zb-h1", 4), ("lockstep", 1)])
def test_lenet_pp2_nccl_matches_single_gpu(tmp_path, free_port, lenet_single, schedule, mb):
    kw = dict(LENET, schedule=schedule, microbatches=mb, p2p="torch")
    _spawn(W.gpu_train_worker, 2, free_port, kw, str(tmp_path))
    res = _load(str(tmp_path), 2)
    assert "Test set:" in res[1]["log"] and res[0]["log"] == ""
    assert all(r["p2p"] == "torch" and not r["graph"] for r in res)
    _close(_merged(res), lenet_single["parts"], 1e-4)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("schedule,mb,graph", [("1f1b", 8, True), ("zb-h1", 4, True), ("lockstep", 1, True),
                                               ("1f1b", 8, False)])
def test_lenet_pp2_symmetric_boundary_matches_single_gpu(tmp_path, free_port, lenet_single, schedule, mb, graph):
    """DEFAULT multi-GPU path: activations / gradients pushed into the neighbour's peer-mapped slots
    (csrc/boundary.cu), the whole step of each rank replayed as one CUDA graph after three eager
    steps (8 steps per epoch: both forms run); evaluation goes through the eval slot."""
    kw = dict(LENET, schedule=schedule, microbatches=mb, use_cuda_graph=graph)
    _spawn(W.gpu_train_worker, 2, free_port, kw, str(tmp_path))
    res = _load(str(tmp_path), 2)
    assert "Test set:" in res[1]["log"] and res[0]["log"] == ""
    assert all(r["p2p"] == "symm" and r["graph"] == graph for r in res)
    assert all(r["graphs"] == (1 if graph else 0) for r in res)
    _close(_merged(res), lenet_single["parts"], 1e-4)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("fused", [True, False])
def test_lenet_dp2_fused_allreduce_sgd_matches_single_gpu(tmp_path, free_port, lenet_single, fused):
    # 2 replicas x batch 30 == one process at batch 60
    kw = dict(LENET, pp=1, dp=2, batch_size=30, fused_collectives=fused)
    _spawn(W.gpu_train_worker, 2, free_port, kw, str(tmp_path))
    res = _load(str(tmp_path), 2)
    assert res[0]["fused_dp"] == fused
    _close(_merged(res), lenet_single["parts"], 1e-4)
    _close(res[0]["parts"], res[1]["parts"], 0.0)      # replicas stay bit-identical


@pytest.mark.skipif(_ngpu() < 4, reason="needs 4 GPUs")
def test_lenet_pp2_dp2_hybrid(tmp_path, free_port, lenet_single):
    kw = dict(LENET, dp=2, batch_size=30, schedule="1f1b", microbatches=2)
    _spawn(W.gpu_train_worker, 4, free_port, kw, str(tmp_path))
    res = _load(str(tmp_path), 4)
    _close(_merged(res), lenet_single["parts"], 1e-4)


MLP = dict(model="mlp", hidden=512, layers=4, classes=256, dtype="bf16", batch_size=512, epochs=1,
           train_samples=3072, test_samples=256, lr=0.02, momentum=0.5, log_interval=2)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("schedule,p2p", [("1f1b", "symm"), ("zb-h1", "symm"), ("zb-h2", "symm"), ("zb-h2", "torch")])
def test_mlp_pp2_matches_single_gpu(tmp_path, free_port, tmp_path_factory, schedule, p2p):
    """symm (default): the boundary GEMMs store their tiles straight into the neighbour's slot and the
    step is one CUDA graph per rank (6 steps: 3 eager + capture + replays); torch: NCCL p2p."""
    d1 = str(tmp_path_factory.mktemp("mlp1"))
    _spawn(W.gpu_train_worker, 1, free_port + 1, dict(MLP, microbatches=4, schedule="gpipe"), d1)
    ref = _load(d1, 1)[0]
    _spawn(W.gpu_train_worker, 2, free_port, dict(MLP, microbatches=4, schedule=schedule, p2p=p2p), str(tmp_path))
    res = _load(str(tmp_path), 2)
    assert all(r["p2p"] == p2p and r["graph"] == (p2p == "symm") for r in res)
    if p2p == "symm":
        assert all(r["direct_writes"] > 0 for r in res), "boundary GEMMs did not write into the peer slots"
    got = {}
    # single GPU holds layers.0-3 in one part; pp2 holds layers.0-1 on each stage
    for r in res:
        base = r["coords"][0] * 2
        for k, v in r["parts"][r["coords"][0]].items():
            l = int(k.split(".")[1])
            got[f"layers.{base + l}.{k.split('.')[2]}"] = v
    for k, v in ref["parts"][0].items():
        torch.testing.assert_close(got[k], v, rtol=2e-2, atol=2e-3, msg=lambda m: f"{k}: {m}")


TPCFG = dict(model="mlp", hidden=512, layers=2, classes=256, dtype="bf16", batch_size=1024, epochs=1,
             train_samples=3072, test_samples=256, lr=0.02, momentum=0.5, log_interval=1, microbatches=2)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("fused", [True, False])
def test_mlp_tp2_matches_single_gpu(tmp_path, free_port, tmp_path_factory, fused):
    """2-way tensor parallel stage (fused AG->GEMM / GEMM->RS kernels, or the NCCL baseline)
    ends with the same weights as the unsharded single-GPU stage."""
    d1 = str(tmp_path_factory.mktemp("tp_ref"))
    _spawn(W.gpu_train_worker, 1, free_port + 1, dict(TPCFG, schedule="gpipe"), d1)
    ref = _load(d1, 1)[0]["parts"][0]
    _spawn(W.gpu_train_worker, 2, free_port, dict(TPCFG, pp=1, tp=2, schedule="gpipe", fused_collectives=fused),
           str(tmp_path))
    res = _load(str(tmp_path), 2)
    shards = {r["coords"][2]: r["parts"][0] for r in res}
    wa = torch.cat([shards[0]["layers.0.weight"], shards[1]["layers.0.weight"]], 0)
    ba = torch.cat([shards[0]["layers.0.bias"], shards[1]["layers.0.bias"]], 0)
    wb = torch.cat([shards[0]["layers.1.weight"], shards[1]["layers.1.weight"]], 1)
    torch.testing.assert_close(wa, ref["layers.0.weight"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(ba, ref["layers.0.bias"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(wb, ref["layers.1.weight"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(shards[0]["layers.1.bias"], ref["layers.1.bias"], rtol=2e-2, atol=2e-3)
    torch.testing.assert_close(shards[0]["layers.1.bias"], shards[1]["layers.1.bias"], rtol=1e-3, atol=1e-4)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("fused", [True, False])
def test_mlp_dp2_fused_wgrad_reduce_scatter_sgd(tmp_path, free_port, tmp_path_factory, fused):
    """2 replicas x batch 512: wgrad GEMM -> reduce-scatter -> sharded SGD -> weight push (fused)
    or NCCL all-reduce + local SGD, both equal to one GPU at batch 1024."""
    cfg = dict(MLP, batch_size=1024, microbatches=2, schedule="gpipe")
    d1 = str(tmp_path_factory.mktemp("dp_ref"))
    _spawn(W.gpu_train_worker, 1, free_port + 1, cfg, d1)
    ref = _load(d1, 1)[0]["parts"][0]
    _spawn(W.gpu_train_worker, 2, free_port, dict(cfg, pp=1, dp=2, batch_size=512, fused_collectives=fused),
           str(tmp_path))
    res = _load(str(tmp_path), 2)
    for k, v in ref.items():
        torch.testing.assert_close(res[0]["parts"][0][k], v, rtol=3e-2, atol=3e-3, msg=lambda m: f"{k}: {m}")
        torch.testing.assert_close(res[0]["parts"][0][k], res[1]["parts"][0][k], rtol=0, atol=0)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_lenet_pp2_native_p2p_engine_matches_single_gpu(tmp_path, free_port, lenet_single):
    """Stage-boundary transfers through csrc/p2p_engine.cu (own NCCL communicators per boundary and
    direction, side streams, event tickets) instead of torch.distributed: same weights."""
    kw = dict(LENET, schedule="1f1b", microbatches=8, p2p="native")
    _spawn(W.gpu_train_worker, 2, free_port, kw, str(tmp_path))
    res = _load(str(tmp_path), 2)
    assert all(r["native_p2p"] for r in res)
    assert "Test set:" in res[1]["log"]
    _close(_merged(res), lenet_single["parts"], 1e-4)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_mlp_dp2_fused_checkpoint_and_resume(tmp_path, free_ports, tmp_path_factory):
    """Trainer.save_checkpoint with the fused DP path (master rows + momentum sharded over the
    replicas): every replica takes part in the gather, replica 0 writes, and a resumed run ends
    where an uninterrupted one does (momentum of EVERY shard restored)."""
    ports = free_ports(3)
    cfg = dict(MLP, batch_size=512, microbatches=2, schedule="gpipe", pp=1, dp=2, train_samples=2048)
    d_full = str(tmp_path_factory.mktemp("full"))
    _spawn(W.gpu_train_worker, 2, ports[0], dict(cfg, epochs=2), d_full)
    ck = str(tmp_path / "ck")
    d1, d2 = str(tmp_path_factory.mktemp("first")), str(tmp_path_factory.mktemp("second"))
    _spawn(W.gpu_train_worker, 2, ports[1], dict(cfg, epochs=1, ckpt_dir=ck), d1)
    assert os.path.exists(os.path.join(ck, "stage0.pt")) and os.path.exists(os.path.join(ck, "stage0.optim.pt"))
    _spawn(W.gpu_train_worker, 2, ports[2], dict(cfg, epochs=2, ckpt_dir=ck, resume=True), d2)
    full, res = _load(d_full, 2)[0]["parts"][0], _load(d2, 2)[0]["parts"][0]
    for k, v in full.items():
        torch.testing.assert_close(res[k], v, rtol=1e-3, atol=1e-4, msg=lambda m: f"{k}: {m}")


@pytest.mark.skipif(_ngpu() < 4, reason="needs 4 GPUs")
@pytest.mark.parametrize("classes", [256, 200])
def test_mlp_pp2_dp2_hybrid_gpu(tmp_path, free_ports, tmp_path_factory, classes):
    """MLP 2 stages x 2 replicas, 4 micro-batches (more than one wgrad group per step: the fused DP
    kernels do not apply, every rank must agree on the NCCL fallback -- with classes=200 the last stage
    is ineligible on its own account too; stages deciding separately used to hang the start-up) == one GPU at the doubled batch."""
    ports = free_ports(2)
    cfg = dict(MLP, classes=classes, batch_size=512, microbatches=4)
    d1 = str(tmp_path_factory.mktemp("ref"))
    _spawn(W.gpu_train_worker, 1, ports[0], dict(cfg, schedule="gpipe"), d1)
    ref = _load(d1, 1)[0]["parts"][0]
    _spawn(W.gpu_train_worker, 4, ports[1], dict(cfg, dp=2, batch_size=256, schedule="zb-h1"), str(tmp_path))
    res = _load(str(tmp_path), 4)
    got = {}
    for r in res:
        if r["coords"][1] != 0:
            continue
        base = r["coords"][0] * 2
        for k, v in r["parts"][r["coords"][0]].items():
            _, l, name = k.split(".")
            got[f"layers.{base + int(l)}.{name}"] = v
    for k, v in ref.items():
        torch.testing.assert_close(got[k], v, rtol=2e-2, atol=2e-3, msg=lambda m: f"{k}: {m}")


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("layout", ["pp2", "dp2"])
def test_stalled_peer_raises_pipeline_timeout_on_gpus(tmp_path, free_port, layout):
    """Failure detection on GPUs: rank 1 stops after 6 steps.  Rank 0's waits are in-kernel spins
    (boundary flag of the symmetric stage boundary / (value, epoch) pairs of the DP all-reduce); each
    carries a globaltimer deadline of --timeout_s, leaves an error record in pinned host memory, and
    the host raises PipelineTimeout naming the peer at its next check."""
    import time
    kw = dict(LENET, train_samples=6000, timeout_s=4.0)
    kw.update(dict(schedule="1f1b", microbatches=4) if layout == "pp2" else dict(pp=1, dp=2, batch_size=30))
    ctx = mp.get_context("spawn")
    d = str(tmp_path)
    procs = [ctx.Process(target=W.stalled_peer_gpu_worker, args=(r, 2, free_port, kw, d)) for r in range(2)]
    for p in procs:
        p.start()
    procs[0].join(120)
    alive0 = procs[0].is_alive()
    for p in procs:
        if p.is_alive():
            p.kill()
    assert not alive0, "rank 0 hung although its peer stalled"
    msg = open(os.path.join(d, "rank0.txt")).read()
    assert msg.startswith("timeout after"), msg
    assert "rank 1" in msg, msg
    assert float(msg.split()[2].rstrip("s:")) < 30.0, msg


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_mlp_interleaved_pp2_v2_matches_single_gpu(tmp_path, free_ports, tmp_path_factory):
    """--virtual_stages 2 on 2 GPUs: 4 virtual stages (chunk c of rank r = stage c * 2 + r), interleaved
    zero-bubble op list, tcgen05 GEMMs per chunk, NCCL p2p across the virtual boundaries (rank 1 -> rank 0
    between the chunks) == one GPU running the same 4-layer MLP."""
    ports = free_ports(2)
    cfg = dict(MLP, microbatches=4)
    d1 = str(tmp_path_factory.mktemp("ref_v"))
    _spawn(W.gpu_train_worker, 1, ports[0], dict(cfg, schedule="gpipe"), d1)
    ref = _load(d1, 1)[0]["parts"][0]
    _spawn(W.gpu_train_worker, 2, ports[1], dict(cfg, virtual_stages=2), str(tmp_path))
    res = _load(str(tmp_path), 2)
    got = {}
    for r in res:
        assert sorted(r["parts"]) == [r["coords"][0], 2 + r["coords"][0]]
        for q, sd in r["parts"].items():
            for k, v in sd.items():
                _, l, name = k.split(".")
                got[f"layers.{q + int(l)}.{name}"] = v          # one layer per virtual stage
    assert sorted(got) == sorted(ref)
    for k, v in ref.items():
        torch.testing.assert_close(got[k], v, rtol=2e-2, atol=2e-3, msg=lambda m: f"{k}: {m}")
