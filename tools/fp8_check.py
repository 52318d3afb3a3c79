"""Timing of the MX-fp8 block-scaled GEMM vs the bf16 tcgen05 GEMM and cuBLAS bf16."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.ops import gemm
from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8


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
for M, N, K in ((8192, 8192, 8192), (2048, 8192, 8192), (4096, 16384, 16384)):
    A = (torch.randn(M, K, device="cuda") * 0.1).bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
    t8 = timeit(lambda: gemm_mxfp8(qa, qb, C), flush)
    t16 = timeit(lambda: gemm.gemm_bf16(A, B, C), flush)
    tq = timeit(lambda: quantize_mxfp8(A), flush)
    fl = 2.0 * M * N * K
    print(f"M={M} N={N} K={K}: mxfp8 {fl / t8 / 1e9:.0f} TF/s ({t8:.3f} ms), bf16 {fl / t16 / 1e9:.0f} TF/s ({t16:.3f} ms), "
          f"quantise A {tq:.3f} ms = {(M * K * 3 + M * K / 32) / tq / 1e6:.0f} GB/s", flush=True)
