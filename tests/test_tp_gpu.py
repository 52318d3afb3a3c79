"""Fused all-gather->GEMM and GEMM->reduce-scatter kernels against NCCL + fp32 matmul
(needs 2 GPUs: `gpurun --gpus 2 -- python -m pytest tests/test_tp_gpu.py -m gpu`)."""
import os

import pytest
import torch
import torch.multiprocessing as mp

from tests import _tp_workers as W

pytestmark = pytest.mark.gpu
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _spawn(fn, world, port, out_dir, timeout=150):
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=fn, args=(r, world, port, out_dir)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    hung = [p for p in procs if p.is_alive()]
    for p in hung:
        p.kill()
    assert not hung, "worker hung"
    assert [p.exitcode for p in procs] == [0] * world
    return [torch.load(os.path.join(out_dir, f"rank{r}.pt")) for r in range(world)]


@pytest.mark.skipif(NGPU < 2, reason="needs 2 GPUs")
def test_all_gather_gemm_matches_nccl_plus_matmul(tmp_path, free_port):
    for res in _spawn(W.ag_gemm_worker, 2, free_port, str(tmp_path)):
        for it in range(3):
            assert res[f"gathered{it}"], "gathered operand differs from ncclAllGather"
            assert res[f"err{it}"] <= 1e-2 * max(1.0, res[f"ref{it}"])


@pytest.mark.skipif(NGPU < 2, reason="needs 2 GPUs")
def test_gemm_reduce_scatter_matches_matmul_plus_nccl(tmp_path, free_port):
    for res in _spawn(W.gemm_rs_worker, 2, free_port, str(tmp_path)):
        for it in range(4):
            assert res[f"err{it}"] <= 1e-2 * max(1.0, res[f"ref{it}"])
