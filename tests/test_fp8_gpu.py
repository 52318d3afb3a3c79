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
