"""MX-fp8 quantisation + block-scaled tcgen05 GEMM vs fp32 oracles (needs a B200)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_quantize_roundtrip_error_bound():
    from simple_distributed_machine_learning_b200.ops.fp8 import quantize_mxfp8
    g = torch.Generator(device="cuda").manual_seed(0)
    x = (torch.randn(300, 512, device="cuda", generator=g) * torch.logspace(-3, 2, 512, device="cuda")).bfloat16()
    q = quantize_mxfp8(x)
    deq = q.dequantize()
    xf = x.float()
    blk = xf.reshape(300, -1, 32)
    amax = blk.abs().amax(dim=2, keepdim=True).clamp_min(1e-30)
    # E4M3 has 3 mantissa bits and the scale costs at most one more bit of the block range
    err = (deq.reshape(300, -1, 32) - blk).abs()
    assert (err <= amax * 2.0 ** -3).all()
    assert ((deq - xf).abs().sum() / xf.abs().sum()).item() < 0.04
    assert (deq[x == 0] == 0).all()


@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (256, 384, 512), (1000, 776, 1024), (2048, 2048, 2048)])
def test_block_scaled_gemm_matches_dequantized_oracle(M, N, K):
    from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * torch.rand(M, 1, device="cuda", generator=g) * 4).bfloat16()
    B = (torch.randn(N, K, device="cuda", generator=g) * 0.3).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
    C = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    gemm_mxfp8(qa, qb, C, bias=bias, relu=True)
    ref = torch.relu(qa.dequantize() @ qb.dequantize().t() + bias)       # exact oracle for the kernel
    scale = max(ref.abs().max().item(), 1.0)
    assert torch.isfinite(C.float()).all()
    assert (C.float() - ref).abs().max().item() <= 1e-2 * scale          # bf16 output rounding
    full = torch.relu(A.float() @ B.float().t() + bias)                 # and close to the bf16 model
    rel = (C.float() - full).norm() / full.norm()
    assert rel.item() < 0.06


def test_block_scaled_gemm_fp32_output_exact():
    from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8
    g = torch.Generator(device="cuda").manual_seed(3)
    A = torch.randn(256, 256, device="cuda", generator=g).bfloat16()
    B = torch.randn(128, 256, device="cuda", generator=g).bfloat16()
    qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
    C = torch.empty(256, 128, device="cuda", dtype=torch.float32)
    gemm_mxfp8(qa, qb, C)
    ref = qa.dequantize().double() @ qb.dequantize().double().t()
    torch.testing.assert_close(C.double(), ref, rtol=1e-4, atol=1e-3)


def test_mlp_training_with_fp8_forward_tracks_bf16():
    from simple_distributed_machine_learning_b200.models.mlp import MLPStage
    from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
    from simple_distributed_machine_learning_b200.parallel.topology import Topology
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1024, 512, generator=g).cuda()
    t = torch.randint(0, 128, (1024,), generator=g).cuda()
    curves = {}
    for dtype in ("bf16", "fp8"):
        st = MLPStage(0, 1, 3, 512, 128, torch.device("cuda"), lr=0.02, momentum=0.5, dtype=dtype)
        eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=2, schedule="gpipe")
        curves[dtype] = [eng.train_step(x, t, 1024).item() for _ in range(25)]
    assert curves["fp8"][-1] < 0.85 * curves["fp8"][0]
    for a, b in zip(curves["fp8"], curves["bf16"]):
        assert a == pytest.approx(b, rel=0.05), (curves["fp8"][::5], curves["bf16"][::5])


def test_dual_quantiser_row_copy_identical_and_transposed_copy_bounded():
    """One pass -> row-wise copy (bit-identical to the single-output quantiser) + TRANSPOSED copy
    scaled along the original rows (the K-major operand of dgrad / wgrad)."""
    from simple_distributed_machine_learning_b200.ops.fp8 import quantize_mxfp8, quantize_mxfp8_dual
    g = torch.Generator(device="cuda").manual_seed(11)
    x = (torch.randn(384, 256, device="cuda", generator=g) * torch.logspace(-2, 2, 256, device="cuda")).bfloat16()
    row, col = quantize_mxfp8_dual(x, want_col=True)
    single = quantize_mxfp8(x)
    assert torch.equal(row.data, single.data)
    assert torch.equal(row.dequantize(), single.dequantize())
    assert col.rows == 256 and col.cols == 384
    deq_t = col.dequantize()                               # [256, 384] = x^T
    xt = x.float().t().contiguous()
    blk = xt.reshape(256, -1, 32)
    amax = blk.abs().amax(dim=2, keepdim=True).clamp_min(1e-30)
    assert ((deq_t.reshape(256, -1, 32) - blk).abs() <= amax * 2.0 ** -3).all()
    assert ((deq_t - xt).abs().sum() / xt.abs().sum()).item() < 0.04
    # transposed copy only
    none, col2 = quantize_mxfp8_dual(x, want_row=False, want_col=True)
    assert none is None and torch.equal(col2.data, col.data) and torch.equal(col2.scales, col.scales)


def test_transposed_buffer_filled_per_microbatch_feeds_wgrad_gemm():
    """Step-wide x^T / dz^T buffers filled one micro-batch at a time; the weight-gradient GEMM
    contracts over a row range of both, with fp32 accumulation into dW across groups."""
    from simple_distributed_machine_learning_b200.ops.fp8 import (MXFP8Transposed, gemm_mxfp8,
                                                                 quantize_mxfp8_dual)
    g = torch.Generator(device="cuda").manual_seed(12)
    B, fi, fo = 512, 256, 384
    x = torch.randn(B, fi, device="cuda", generator=g).bfloat16()
    dz = (torch.randn(B, fo, device="cuda", generator=g) * 0.01).bfloat16()
    xT, dzT = MXFP8Transposed(B, fi, "cuda"), MXFP8Transposed(B, fo, "cuda")
    for lo in range(0, B, 128):
        quantize_mxfp8_dual(x[lo:lo + 128], want_row=False, col_into=xT, row_offset=lo)
        quantize_mxfp8_dual(dz[lo:lo + 128], want_row=False, col_into=dzT, row_offset=lo)
    _, whole = quantize_mxfp8_dual(x, want_row=False, want_col=True)
    assert torch.equal(xT.data, whole.data) and torch.equal(xT.scales, whole.scales)
    dW = torch.empty(fo, fi, device="cuda", dtype=torch.float32)
    gemm_mxfp8(dzT.rows_slice(0, 256), xT.rows_slice(0, 256), dW)
    gemm_mxfp8(dzT.rows_slice(256, 512), xT.rows_slice(256, 512), dW, accumulate=True)
    exact = dzT.rows_slice(0, 512).dequantize().double() @ xT.rows_slice(0, 512).dequantize().double().t()
    torch.testing.assert_close(dW.double(), exact, rtol=1e-4, atol=1e-5)
    full = dz.float().t() @ x.float()
    assert ((dW - full).norm() / full.norm()).item() < 0.06


def test_fp8_dgrad_matches_bf16_dgrad():
    from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8_dual
    g = torch.Generator(device="cuda").manual_seed(13)
    n, fi, fo = 256, 384, 512
    W = (torch.randn(fo, fi, device="cuda", generator=g) * 0.05).bfloat16()
    dz = (torch.randn(n, fo, device="cuda", generator=g) * 0.01).bfloat16()
    xl = torch.randn(n, fi, device="cuda", generator=g).bfloat16()          # ReLU mask source
    _, wT = quantize_mxfp8_dual(W, want_row=False, want_col=True)          # W^T [fi, fo], K = fo
    dzq, _ = quantize_mxfp8_dual(dz)
    dx = torch.empty(n, fi, device="cuda", dtype=torch.bfloat16)
    gemm_mxfp8(dzq, wT, dx, aux=xl)
    ref = (dz.float() @ W.float()) * (xl.float() > 0)
    assert ((dx.float() - ref).norm() / ref.norm()).item() < 0.06


def test_mlp_training_with_fp8_backward_tracks_bf16(monkeypatch):
    """dtype fp8 with SDML_FP8_BWD=1: forward, dgrad and wgrad all on the block-scaled kernels."""
    from simple_distributed_machine_learning_b200.models.mlp import MLPStage
    from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
    from simple_distributed_machine_learning_b200.parallel.topology import Topology
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1024, 512, generator=g).cuda()
    t = torch.randint(0, 128, (1024,), generator=g).cuda()
    curves, grads = {}, {}
    for dtype, bwd in (("bf16", "0"), ("fp8", "1")):
        monkeypatch.setenv("SDML_FP8_BWD", bwd)
        st = MLPStage(0, 1, 3, 512, 128, torch.device("cuda"), lr=0.02, momentum=0.5, dtype=dtype)
        st.w_group = 1                      # one wgrad GEMM per micro-batch: exercises accumulate
        eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=4, schedule="gpipe")
        curves[dtype] = [eng.train_step(x, t, 1024).item() for _ in range(25)]
        if dtype == "fp8":
            assert st.fp8_backward and all(st._fp8_bwd_ok)
    assert curves["fp8"][-1] < 0.85 * curves["fp8"][0]
    for a, b in zip(curves["fp8"], curves["bf16"]):
        assert a == pytest.approx(b, rel=0.06), (curves["fp8"][::5], curves["bf16"][::5])


@pytest.mark.parametrize("tile", ["1cta", "2cta128", "2cta192"])
@pytest.mark.parametrize("M,N,K", [(256, 192, 256), (512, 576, 1024), (1000, 776, 1024)])
def test_block_scaled_gemm_tile_variants_agree_with_oracle(tile, M, N, K):
    """Single-CTA 128x128 kernel and the CTA-pair (cta_group::2) 256x128 / 256x192 kernels, incl.
    tiles that start mid scale-atom (odd 192-wide tiles) and ragged M / N edges."""
    from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8
    g = torch.Generator(device="cuda").manual_seed(M * 3 + N * 5 + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * torch.rand(M, 1, device="cuda", generator=g) * 4).bfloat16()
    B = (torch.randn(N, K, device="cuda", generator=g) * 0.3).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
    C = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    gemm_mxfp8(qa, qb, C, bias=bias, relu=True, tile=tile)
    ref = torch.relu(qa.dequantize() @ qb.dequantize().t() + bias)
    assert torch.isfinite(C.float()).all()
    assert (C.float() - ref).abs().max().item() <= 1e-2 * max(ref.abs().max().item(), 1.0)
    C32 = torch.ones(M, N, device="cuda", dtype=torch.float32)
    gemm_mxfp8(qa, qb, C32, accumulate=True, tile=tile)
    ref32 = 1.0 + qa.dequantize().double() @ qb.dequantize().double().t()
    torch.testing.assert_close(C32.double(), ref32, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("M,N,K", [(256, 384, 256), (512, 768, 1024), (2048, 2048, 512)])
def test_fused_output_quantisation_is_bit_identical_to_the_quantiser_pass(M, N, K, monkeypatch):
    """sdml_gemm_mxfp8_2cta_qout: the epilogue's row-wise and transposed fp8 copies of C must equal
    sdml_quant_mxfp8_dual(C) byte for byte (payload and scale atoms)."""
    from simple_distributed_machine_learning_b200.ops import fp8 as F
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * torch.rand(M, 1, device="cuda", generator=g) * 4).bfloat16()
    B = (torch.randn(N, K, device="cuda", generator=g) * 0.3).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    qa, qb = F.quantize_mxfp8(A), F.quantize_mxfp8(B)
    outs = {}
    for fused in (False, True):
        monkeypatch.setattr(F, "_FUSED_QOUT", fused)
        C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        row = F.MXFP8Tensor(torch.zeros(M, N, dtype=torch.uint8, device="cuda"),
                            torch.zeros(F.sf_bytes(M, N), dtype=torch.uint8, device="cuda"), M, N)
        T = F.MXFP8Transposed(2 * M, N, "cuda")
        F.gemm_mxfp8(qa, qb, C, bias=bias, relu=True, q_row_out=row, q_col_into=T, q_row_offset=M)
        outs[fused] = (C, row, T)
    assert torch.equal(outs[True][0], outs[False][0])
    assert torch.equal(outs[True][1].data, outs[False][1].data)
    assert torch.equal(outs[True][1].scales, outs[False][1].scales)
    assert torch.equal(outs[True][2].data[:, M:], outs[False][2].data[:, M:])
    k0 = (M // 128) * outs[True][2].groups * 512
    assert torch.equal(outs[True][2].scales[k0:], outs[False][2].scales[k0:])


def test_tma_store_epilogue_matches_direct_stores():
    """The CTA-pair kernel's two bf16 epilogues (per-thread stores for local outputs, 32 x 64 TMA boxes
    for a pipeline neighbour's slot) must write the same bits; SDML_FP8_TMA_STORE forces one form, so
    each runs in its own process."""
    import subprocess
    import sys
    code = r"""
import sys, torch
sys.path.insert(0, %r)
from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8
out = {}
for (M, N, K) in [(256, 192, 256), (1000, 776, 1024), (2048, 8192, 1024)]:
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * 2).bfloat16()
    B = (torch.randn(N, K, device="cuda", generator=g) * 0.3).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    aux = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    for tile in ("2cta192", "2cta128"):
        C = torch.full((M + 3, N), 7.0, device="cuda", dtype=torch.bfloat16)     # guard rows past M
        gemm_mxfp8(quantize_mxfp8(A), quantize_mxfp8(B), C[:M], bias=bias, relu=True, tile=tile)
        D = torch.full((M, N), 7.0, device="cuda", dtype=torch.bfloat16)
        gemm_mxfp8(quantize_mxfp8(A), quantize_mxfp8(B), D, aux=aux, tile=tile)
        assert (C[M:] == 7.0).all(), "wrote past the last row"
        out[(M, N, K, tile)] = (C[:M].cpu(), D.cpu())
torch.save(out, sys.argv[1])
""" % str(__import__("pathlib").Path(__file__).resolve().parent.parent)
    import os
    import tempfile
    res = []
    with tempfile.TemporaryDirectory() as d:
        for mode in ("0", "1"):
            path = os.path.join(d, f"m{mode}.pt")
            env = dict(os.environ, SDML_FP8_TMA_STORE=mode)
            subprocess.run([sys.executable, "-c", code, path], check=True, env=env, timeout=300)
            res.append(torch.load(path))
    assert sorted(res[0]) == sorted(res[1])
    for k in res[0]:
        for a, b in zip(res[0][k], res[1][k]):
            assert torch.equal(a, b), f"epilogues differ for {k}"
