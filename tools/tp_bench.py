"""Fused tensor-parallel kernels vs the NCCL + cuBLAS baseline (run under torchrun, 2+ GPUs).
Device-timed with CUDA events, max over ranks, L2 flushed between iterations."""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.parallel.tp_ops import AllGatherGemm, GemmReduceScatter  # noqa: E402


def timeit(fn, iters=8, flush=None):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    ctrl = dist.new_group(list(range(world)), backend="gloo")
    M = int(os.environ.get("TP_M", 4096))
    H = int(os.environ.get("TP_H", 16384))
    hs = H // world
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    peaks = json.load(open(Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json")) \
        if (Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").exists() else {"bf16_tflops": 1590.0}
    peak = peaks["bf16_tflops"] * 1e12
    link = 770e9
    out = {"world": world, "M": M, "H": H}

    # ---- all-gather -> GEMM:  [M, H] (row-sharded) x Wa_t [H/tp, H]^T ----
    ag = AllGatherGemm(list(range(world)), rank, ctrl, M, H, dev)
    Wa = (torch.randn(hs, H, device=dev) * 0.02).bfloat16()
    shard = (torch.randn(M // world, H, device=dev) * 0.5).bfloat16()
    ag.local_rows().copy_(shard)
    C = torch.empty(M, hs, device=dev, dtype=torch.bfloat16)
    full = torch.empty(M, H, device=dev, dtype=torch.bfloat16)

    def fused_ag():
        ag.run(Wa, C, relu=True)

    def nccl_ag():
        dist.all_gather_into_tensor(full, shard)
        torch.matmul(full, Wa.t(), out=C)
        torch.relu_(C)

    flops = 2.0 * M * H * hs
    t_f, t_n = timeit(fused_ag, flush=flush), timeit(nccl_ag, flush=flush)
    bound = max(flops / peak, (world - 1) / world * M * H * 2 / link)
    out["ag_gemm"] = {"fused_ms": t_f, "nccl_cublas_ms": t_n, "roofline_ms": bound * 1e3,
                      "fused_frac_of_roofline": bound * 1e3 / t_f, "speedup_vs_nccl": t_n / t_f,
                      "tflops": flops / t_f / 1e9}

    # ---- GEMM -> reduce-scatter:  y1_t [M, H/tp] x Wb_t [H, H/tp]^T ----
    rs = GemmReduceScatter(list(range(world)), rank, ctrl, M, H, dev)
    A = (torch.randn(M, hs, device=dev) * 0.5).bfloat16()
    Wb = (torch.randn(H, hs, device=dev) * 0.02).bfloat16()
    bias = torch.randn(H, device=dev)
    outb = torch.empty(M // world, H, device=dev, dtype=torch.bfloat16)
    part = torch.empty(M, H, device=dev, dtype=torch.bfloat16)
    red = torch.empty(M // world, H, device=dev, dtype=torch.bfloat16)

    def fused_rs():
        par = rs.run(A, Wb)
        rs.finalize(par, outb, bias=bias, relu=True)

    def nccl_rs():
        torch.matmul(A, Wb.t(), out=part)
        dist.reduce_scatter_tensor(red, part)
        torch.relu_(red.add_(bias.to(torch.bfloat16)))

    t_f, t_n = timeit(fused_rs, flush=flush), timeit(nccl_rs, flush=flush)
    bound = max(flops / peak, (world - 1) / world * M * H * 4 / link)
    out["gemm_rs"] = {"fused_ms": t_f, "nccl_cublas_ms": t_n, "roofline_ms": bound * 1e3,
                      "fused_frac_of_roofline": bound * 1e3 / t_f, "speedup_vs_nccl": t_n / t_f,
                      "tflops": flops / t_f / 1e9,
                      "note": "baseline reduce-scatters bf16 partials; the fused path adds fp32 partials"}
    if rank == 0:
        print(json.dumps(out, indent=1))
        Path("gpurun_out").mkdir(exist_ok=True)
        Path(f"gpurun_out/tp_bench_n{world}.json").write_text(json.dumps(out, indent=1))
    dist.barrier()
    dist.destroy_process_group()


main()
