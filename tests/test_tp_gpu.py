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


@pytest.mark.skipif(NGPU < 2, reason="needs 2 GPUs")
def test_fp8_gemm_reduce_scatter_matches_dequantised_matmul_plus_nccl(tmp_path, free_port):
    """Block-scaled fp8 GEMM whose epilogue scatters the bf16 partial rows into the owners' landing slots
    (sdml_gemm_mxfp8_2cta_rs) + sdml_rs_finalize == fp32 product of the dequantised operands reduced by NCCL;
    four calls in a row exercise both landing parities and the counters."""
    for res in _spawn(W.fp8_gemm_rs_worker, 2, free_port, str(tmp_path)):
        for it in range(4):
            assert res[f"err{it}"] <= 1.5e-2 * max(1.0, res[f"ref{it}"])


@pytest.mark.skipif(NGPU < 2, reason="needs 2 GPUs")
def test_fp8_push_all_gather_is_bit_identical_to_gather_then_quantise(tmp_path, free_port):
    """tp_ops.Fp8AllGather: every rank quantises its shard and pushes fp8 rows + scale atoms into its
    peer's buffers; the result must equal quantising the NCCL-gathered operand, byte for byte (row-wise and
    transposed copies, 2 slots x 3 steps), and feed the block-scaled GEMM correctly."""
    for res in _spawn(W.fp8_all_gather_worker, 2, free_port, str(tmp_path)):
        assert res["row"] and res["row_sf"], "row-wise operand differs"
        assert res["col"] and res["col_sf"], "transposed operand differs"
        assert res["gemm_err"] <= 1e-2 * res["gemm_ref"]


def test_tp_stage_eval_forward_matches_mlp_stage():
    """The TP stage's evaluation pass on the GPU kernels (bf16 GEMM + fp32-output partial GEMM);
    with tp=1 it must score a batch exactly like the unsharded MLP stage from the same seed."""
    from simple_distributed_machine_learning_b200.models.mlp import MLPStage
    from simple_distributed_machine_learning_b200.parallel.tensor_parallel import TPMLPStage
    dev = torch.device("cuda")
    H, C, n = 512, 256, 200                      # ragged batch on purpose
    x = torch.randn(n, H, generator=torch.Generator().manual_seed(3)).to(dev)
    t = torch.randint(0, C, (n,), generator=torch.Generator().manual_seed(4)).to(dev)
    a = MLPStage(0, 1, 2, H, C, dev, seed=5)
    b = TPMLPStage(0, 1, 2, H, C, dev, seed=5, tp=1, tp_idx=0, backend="nccl")
    for st in (a, b):
        st.eval_begin()
        st.eval_forward(x, t)
    torch.cuda.synchronize()
    assert b.loss_sum.item() == pytest.approx(a.loss_sum.item(), rel=1e-3)
    assert abs(int(b.correct.item()) - int(a.correct.item())) <= 1
    ref = torch.nn.functional.cross_entropy(
        (torch.relu(x.bfloat16().float() @ a.w_master[0].bfloat16().float().t() + a.b_master[0]).bfloat16().float()
         @ a.w_master[1].bfloat16().float().t() + a.b_master[1]), t, reduction="sum")
    assert b.loss_sum.item() == pytest.approx(ref.item(), rel=2e-2)
