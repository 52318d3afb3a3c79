"""Stand-alone correctness + timing sweep of the tcgen05 GEMM (run on a B200 via gpurun).

Prints one line per case: max abs error against an fp32 PyTorch oracle, and for the
timing cases TFLOP/s with CUDA events (L2 flushed between iterations) next to
torch.matmul (cuBLAS) on the same shapes.
"""
from __future__ import annotations

import json
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.ops import gemm  # noqa: E402


def oracle(A, B, a_mn, b_mn, bias, aux, relu, alpha, acc_into):
    a = A.float().t() if a_mn else A.float()
    b = B.float().t() if b_mn else B.float()
    c = alpha * (a @ b.t())
    if bias is not None:
        c = c + bias
    if relu:
        c = torch.relu(c)
    if aux is not None:
        c = c * (aux.float() > 0)
    if acc_into is not None:
        c = c + acc_into
    return c


def run_case(M, N, K, a_mn, b_mn, bias=False, aux=False, relu=False, out_fp32=False, accumulate=False,
             alpha=1.0, bn=0):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    A = (torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g) * 0.5).bfloat16()
    B = (torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g) * 0.5).bfloat16()
    bias_t = torch.randn(N, device="cuda", generator=g) if bias else None
    aux_t = torch.randn(M, N, device="cuda", generator=g).bfloat16() if aux else None
    if accumulate:
        C = torch.randn(M, N, device="cuda", generator=g)
        acc0 = C.clone()
    else:
        C = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float32 if out_fp32 else torch.bfloat16)
        acc0 = None
    gemm.gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bias=bias_t, aux=aux_t, relu=relu,
                   accumulate=accumulate, alpha=alpha, bn=bn)
    torch.cuda.synchronize()
    ref = oracle(A, B, a_mn, b_mn, bias_t, aux_t, relu, alpha, acc0)
    err = (C.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    tol = (2e-2 if not out_fp32 else 2e-3) * max(scale, 1.0)
    ok = bool(err <= tol) and bool(torch.isfinite(C.float()).all())
    print(f"case M={M} N={N} K={K} a_mn={a_mn} b_mn={b_mn} bias={bias} aux={aux} relu={relu} "
          f"fp32={out_fp32} acc={accumulate} bn={bn}: max_err={err:.4g} ref_max={scale:.4g} {'OK' if ok else 'FAIL'}",
          flush=True)
    return ok


def time_case(M, N, K, a_mn, b_mn, bn=0, iters=10):
    A = (torch.randn((K, M) if a_mn else (M, K), device="cuda") * 0.1).bfloat16()
    B = (torch.randn((K, N) if b_mn else (N, K), device="cuda") * 0.1).bfloat16()
    C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def ours():
        gemm.gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bn=bn)

    a_l = A.t() if a_mn else A
    b_l = B if b_mn else B.t()

    def cublas():
        torch.matmul(a_l, b_l, out=C)

    res = {}
    for name, fn in (("ours", ours), ("cublas", cublas)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        times = []
        for _ in range(iters):
            flush.zero_()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        times.sort()
        med = times[len(times) // 2]
        res[name] = {"ms_med": med, "ms_min": times[0], "tflops_med": 2.0 * M * N * K / med / 1e9,
                     "tflops_best": 2.0 * M * N * K / times[0] / 1e9}
    print(f"time M={M} N={N} K={K} a_mn={a_mn} b_mn={b_mn} bn={bn}: "
          f"ours {res['ours']['tflops_med']:.0f} TF/s (best {res['ours']['tflops_best']:.0f}), "
          f"cublas {res['cublas']['tflops_med']:.0f} TF/s (best {res['cublas']['tflops_best']:.0f})", flush=True)
    return res


def main():
    quick = "--quick" in sys.argv
    ok = True
    t0 = time.time()
    # single tile, K-major both: the simplest path first
    ok &= run_case(128, 64, 64, 0, 0, bn=64)
    ok &= run_case(128, 128, 128, 0, 0, bn=128)
    ok &= run_case(128, 256, 256, 0, 0, bn=256)
    # majors
    for a_mn, b_mn in ((0, 1), (1, 0), (1, 1)):
        ok &= run_case(128, 128, 128, a_mn, b_mn, bn=128)
        ok &= run_case(256, 512, 320, a_mn, b_mn)
    # multi-tile persistent + ragged edges
    ok &= run_case(1000, 776, 520, 0, 0)
    ok &= run_case(1000, 776, 520, 0, 1)
    ok &= run_case(1000, 776, 520, 1, 1, out_fp32=True)
    ok &= run_case(4096, 4096, 1024, 0, 0)
    # epilogues
    ok &= run_case(512, 512, 256, 0, 0, bias=True, relu=True)
    ok &= run_case(512, 512, 256, 0, 1, aux=True)
    ok &= run_case(512, 512, 256, 1, 1, out_fp32=True, accumulate=True, alpha=0.5)
    ok &= run_case(60, 56, 320, 0, 0, bias=True, relu=True)
    # CTA-pair kernels (cta_group::2): 256 x 256 and 256 x 128 tiles, every operand major
    for bn in (1256, 1128):
        for a_mn, b_mn in ((0, 0), (0, 1), (1, 0), (1, 1)):
            ok &= run_case(256, 256, 128, a_mn, b_mn, bn=bn)
            ok &= run_case(1000, 776, 520, a_mn, b_mn, bn=bn, out_fp32=bool(a_mn and b_mn))
    ok &= run_case(2048, 2048, 2048, 0, 0, bias=True, relu=True, bn=1256)
    ok &= run_case(2048, 1024, 512, 1, 1, out_fp32=True, accumulate=True, alpha=0.5, bn=1256)
    print(f"correctness {'PASS' if ok else 'FAIL'} in {time.time() - t0:.1f}s", flush=True)
    out = {"ok": ok}
    if ok and not quick:
        out["fwd_8192"] = time_case(8192, 8192, 8192, 0, 0)
        out["fwd_2048x8192x8192"] = time_case(2048, 8192, 8192, 0, 0)
        out["dgrad_2048x8192x8192"] = time_case(2048, 8192, 8192, 0, 1)
        out["wgrad_8192x8192x2048"] = time_case(8192, 8192, 2048, 1, 1)
        out["fwd_8192_1cta256"] = time_case(8192, 8192, 8192, 0, 0, bn=256)
        out["fwd_8192_2cta128"] = time_case(8192, 8192, 8192, 0, 0, bn=1128)
        out["fwd_4736x8192x8192"] = time_case(4736, 8192, 8192, 0, 0)
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/gemm_check.json").write_text(json.dumps(out, indent=1))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
