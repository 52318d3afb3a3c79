"""CTA-pair MX-fp8 GEMM: numerics against the dequantised oracle / the single-CTA kernel, then timing."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8

tiles = sys.argv[1].split(",") if len(sys.argv) > 1 else ["2cta128", "2cta192"]
torch.manual_seed(0)
ok = True
for tile in tiles:
    for (M, N, K) in ((256, 128, 128), (256, 192, 256), (256, 384, 512), (512, 576, 1024), (1000, 776, 1024), (2048, 2048, 2048)):
        A = (torch.randn(M, K, device="cuda") * torch.rand(M, 1, device="cuda") * 4).bfloat16()
        B = (torch.randn(N, K, device="cuda") * 0.3).bfloat16()
        bias = torch.randn(N, device="cuda")
        qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
        ref = torch.relu(qa.dequantize() @ qb.dequantize().t() + bias)
        C = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
        gemm_mxfp8(qa, qb, C, bias=bias, relu=True, tile=tile)
        torch.cuda.synchronize()
        err = (C.float() - ref).abs().max().item()
        scale = max(ref.abs().max().item(), 1.0)
        good = bool(torch.isfinite(C.float()).all()) and err <= 1e-2 * scale
        # fp32 accumulate
        C32 = torch.ones(M, N, device="cuda", dtype=torch.float32)
        gemm_mxfp8(qa, qb, C32, accumulate=True, tile=tile)
        ref32 = 1.0 + qa.dequantize().double() @ qb.dequantize().double().t()
        err32 = (C32.double() - ref32).abs().max().item()
        good32 = err32 <= 1e-3 * max(ref32.abs().max().item(), 1.0)
        ok &= good and good32
        print(f"{tile} M={M} N={N} K={K}: max err {err:.4f} (scale {scale:.1f}) {'ok' if good else 'FAIL'}; "
              f"fp32+acc err {err32:.2e} {'ok' if good32 else 'FAIL'}", flush=True)
        if not good:
            bad = ((C.float() - ref).abs() > 1e-2 * scale) | ~torch.isfinite(C.float())
            rows = bad.any(1).nonzero().flatten()[:8].tolist()
            cols = bad.any(0).nonzero().flatten()
            print(f"   bad rows (first) {rows}  bad cols range {cols.min().item()}..{cols.max().item()} n={cols.numel()}")
print("NUMERICS", "PASS" if ok else "FAIL", flush=True)


def timeit(fn, flush, iters=8):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for M, N, K in ((8192, 8192, 8192), (2048, 8192, 8192), (8192, 8192, 16384), (4096, 16384, 16384)):
    A = (torch.randn(M, K, device="cuda") * 0.1).bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
    fl = 2.0 * M * N * K
    line = f"M={M} N={N} K={K}:"
    for tile in ["1cta"] + tiles:
        t = timeit(lambda: gemm_mxfp8(qa, qb, C, tile=tile), flush)
        line += f"  {tile} {fl / t / 1e9:.0f} TF/s ({t:.3f} ms)"
    print(line, flush=True)
