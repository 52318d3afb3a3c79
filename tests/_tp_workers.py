"""2-GPU workers for the fused tensor-parallel kernels (spawned by tests/test_tp_gpu.py)."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch
import torch.distributed as dist


def _init(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    return dist.new_group(list(range(world)), backend="gloo")


def ag_gemm_worker(rank, world, port, out_dir):
    from simple_distributed_machine_learning_b200.parallel.tp_ops import AllGatherGemm
    ctrl = _init(rank, world, port)
    dev = torch.device("cuda", rank)
    M, K, N = 2048, 1024, 1536
    res = {}
    op = AllGatherGemm(list(range(world)), rank, ctrl, M, K, dev)
    for it, b_mn in enumerate((0, 1, 0)):
        g = torch.Generator(device=dev).manual_seed(100 * it + rank)
        shard = (torch.randn(M // world, K, device=dev, generator=g) * 0.5).bfloat16()
        gw = torch.Generator(device=dev).manual_seed(7 + it)
        W = (torch.randn((K, N) if b_mn else (N, K), device=dev, generator=gw) * 0.1).bfloat16()
        bias = torch.randn(N, device=dev, generator=gw)
        op.local_rows().copy_(shard)
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        op.run(W, C, b_mn=b_mn, bias=bias, relu=True)
        # NCCL + cuBLAS baseline
        full = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
        dist.all_gather_into_tensor(full, shard)
        ref = torch.relu(full.float() @ (W.float() if b_mn else W.float().t()) + bias)
        torch.cuda.synchronize()
        res[f"err{it}"] = (C.float() - ref).abs().max().item()
        res[f"ref{it}"] = ref.abs().max().item()
        res[f"gathered{it}"] = bool(torch.equal(op.full, full))
    torch.save(res, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def gemm_rs_worker(rank, world, port, out_dir):
    from simple_distributed_machine_learning_b200.parallel.tp_ops import GemmReduceScatter
    ctrl = _init(rank, world, port)
    dev = torch.device("cuda", rank)
    M, K, N = 2048, 1024, 1536
    res = {}
    op = GemmReduceScatter(list(range(world)), rank, ctrl, M, N, dev)
    for it, b_mn in enumerate((0, 1, 0, 0)):
        g = torch.Generator(device=dev).manual_seed(100 * it + rank)
        A = (torch.randn(M, K, device=dev, generator=g) * 0.5).bfloat16()
        W = (torch.randn((K, N) if b_mn else (N, K), device=dev, generator=g) * 0.1).bfloat16()
        gb = torch.Generator(device=dev).manual_seed(5 + it)
        bias = torch.randn(N, device=dev, generator=gb)
        par = op.run(A, W, b_mn=b_mn)
        out = torch.empty(M // world, N, device=dev, dtype=torch.bfloat16)
        op.finalize(par, out, bias=bias, relu=True)
        partial = A.float() @ (W.float() if b_mn else W.float().t())
        red = torch.empty(M // world, N, device=dev)
        dist.reduce_scatter_tensor(red, partial)
        ref = torch.relu(red + bias)
        torch.cuda.synchronize()
        res[f"err{it}"] = (out.float() - ref).abs().max().item()
        res[f"ref{it}"] = ref.abs().max().item()
    torch.save(res, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def fp8_gemm_rs_worker(rank, world, port, out_dir):
    """MX-fp8 GEMM -> reduce-scatter against the dequantised operands' fp32 product + ncclReduceScatter."""
    from simple_distributed_machine_learning_b200.ops.fp8 import quantize_mxfp8
    from simple_distributed_machine_learning_b200.parallel.tp_ops import Fp8GemmReduceScatter
    ctrl = _init(rank, world, port)
    dev = torch.device("cuda", rank)
    M, K, N = 2048, 1024, 1536
    res = {}
    op = Fp8GemmReduceScatter(list(range(world)), rank, ctrl, M, N, dev)
    for it in range(4):
        g = torch.Generator(device=dev).manual_seed(100 * it + rank)
        A = (torch.randn(M, K, device=dev, generator=g) * 0.5).bfloat16()
        W = (torch.randn(N, K, device=dev, generator=g) * 0.1).bfloat16()
        gb = torch.Generator(device=dev).manual_seed(5 + it)
        bias = torch.randn(N, device=dev, generator=gb)
        aux = torch.randn(M // world, N, device=dev, generator=gb).bfloat16() if it == 2 else None
        qa, qw = quantize_mxfp8(A), quantize_mxfp8(W)
        par = op.run(qa, qw)
        out = torch.empty(M // world, N, device=dev, dtype=torch.bfloat16)
        op.finalize(par, out, bias=bias, aux=aux, relu=aux is None)
        partial = qa.dequantize() @ qw.dequantize().t()
        red = torch.empty(M // world, N, device=dev)
        dist.reduce_scatter_tensor(red, partial)
        ref = torch.relu(red + bias) if aux is None else (red + bias) * (aux.float() > 0)
        torch.cuda.synchronize()
        res[f"err{it}"] = (out.float() - ref).abs().max().item()
        res[f"ref{it}"] = ref.abs().max().item()
    torch.save(res, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def fp8_all_gather_worker(rank, world, port, out_dir):
    """Push-based fp8 all-gather: after gather() every rank's operand buffers must hold, bit for bit, what
    quantising the NCCL-gathered bf16 operand as a whole would produce (row-wise copy + scale atoms, and the
    step-wide transposed copy + its atoms), for several slots and several steps (slot reuse + acknowledgements)."""
    from simple_distributed_machine_learning_b200.ops.fp8 import MXFP8Transposed, gemm_mxfp8, quantize_mxfp8, quantize_mxfp8_dual
    from simple_distributed_machine_learning_b200.parallel.tp_ops import Fp8AllGather
    ctrl = _init(rank, world, port)
    dev = torch.device("cuda", rank)
    M, K, N, slots = 1024, 512, 768, 2
    op = Fp8AllGather(list(range(world)), rank, ctrl, M, K, dev, slots=slots, timeout_s=20.0)
    res = {"row": True, "row_sf": True, "col": True, "col_sf": True, "gemm_err": 0.0, "gemm_ref": 1.0}
    W = (torch.randn(N, K, device=dev, generator=torch.Generator(device=dev).manual_seed(3)) * 0.1).bfloat16()
    qw = quantize_mxfp8(W)
    for step in range(3):
        ref_t = MXFP8Transposed(slots * M, K, dev)
        for slot in range(slots):
            g = torch.Generator(device=dev).manual_seed(1000 * step + 10 * slot + rank)
            shard = (torch.randn(M // world, K, device=dev, generator=g) * (0.5 + rank)).bfloat16()
            q = op.gather(slot, shard)
            full = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
            dist.all_gather_into_tensor(full, shard)
            ref_row, _ = quantize_mxfp8_dual(full, col_into=ref_t, row_offset=slot * M)
            C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            gemm_mxfp8(q, qw, C)
            torch.cuda.synchronize()
            res["row"] &= bool(torch.equal(q.data, ref_row.data))
            res["row_sf"] &= bool(torch.equal(q.scales, ref_row.scales))
            want = ref_row.dequantize() @ qw.dequantize().t()
            res["gemm_err"] = max(res["gemm_err"], (C.float() - want).abs().max().item())
            res["gemm_ref"] = max(res["gemm_ref"], want.abs().max().item())
        torch.cuda.synchronize()
        res["col"] &= bool(torch.equal(op.transposed.data, ref_t.data))
        res["col_sf"] &= bool(torch.equal(op.transposed.scales, ref_t.scales))
        op.end_step()
        op.check()
    torch.cuda.synchronize()
    torch.save(res, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()
