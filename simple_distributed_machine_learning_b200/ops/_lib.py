"""(Loader for the native libraries; the reference has no native code -- /root/reference is one
Python script -- so this file has no counterpart there.)

ctypes loader for the native libraries.

The CUDA library is mandatory on a GPU box: if CUDA is available and the library is
missing or fails to load we raise instead of silently falling back to eager PyTorch
(the driver records which ``.so`` files were loaded).
"""
from __future__ import annotations

import ctypes
import threading
from ctypes import c_char_p, c_float, c_int, c_longlong, c_void_p
from pathlib import Path

from . import _build

_lock = threading.Lock()
_cuda_lib: ctypes.CDLL | None = None
_sched_lib: ctypes.CDLL | None = None


class NativeError(RuntimeError):
    pass


def _declare_cuda(lib: ctypes.CDLL) -> None:
    P, I, L, F = c_void_p, c_int, c_longlong, c_float
    lib.sdml_last_error.restype = c_char_p
    lib.sdml_last_error.argtypes = []
    lib.sdml_launch_count.restype = c_longlong
    lib.sdml_launch_count.argtypes = []
    lib.sdml_abi_version.restype = c_int
    if hasattr(lib, "sdml_set_wait_timeout"):
        lib.sdml_set_wait_timeout.restype = c_int
        lib.sdml_set_wait_timeout.argtypes = [ctypes.c_ulonglong, c_void_p]
    sigs = {
        # A, B, C, M, N, K, lda, ldb, ldc, a_mn, b_mn, bias, aux, ld_aux, relu, accumulate,
        # out_fp32, alpha, bn, stream
        "sdml_gemm_bf16": [P, P, P, I, I, I, L, L, L, I, I, P, P, L, I, I, I, F, I, P],
        # logits, ld, dtype, targets, rows, classes, loss_sum, dlogits, ld_d, grad_scale, correct, stream
        "sdml_softmax_ce": [P, L, I, P, I, I, P, P, L, F, P, P],
        # n_tensors, params**, grads**, moms**, lowp**, sizes*, lr, momentum, weight_decay, grad_scale, first_step, zero_grad, stream
        "sdml_sgd_multi": [I, P, P, P, P, P, F, F, F, F, I, I, P],
        # n_tensors, params**, grads**, m**, v**, lowp**, sizes*, lr, b1, b2, eps, wd, grad_scale, step, zero_grad, stream
        "sdml_adam_multi": [I, P, P, P, P, P, P, F, F, F, F, F, F, I, I, P],
        # x, w, b, y, M, N, K, relu, drop_p, seed, offset, offset_dev, stream
        "sdml_linear_f32_fwd": [P, P, P, P, I, I, I, I, F, L, L, P, P],
        # x, w, y, dy, dx, dw, db, M, N, K, relu, drop_p, seed, offset, offset_dev, accumulate, stream
        "sdml_linear_f32_bwd": [P, P, P, P, P, P, P, I, I, I, I, F, L, L, P, I, P],
        # x(bf16), ld, rows, cols, out(fp32), accumulate, stream
        "sdml_colsum_bf16": [P, L, I, I, P, I, P],
        # src(fp32), dst(bf16), n, stream
        "sdml_cast_f32_bf16": [P, P, L, P],
        # symmetric memory (csrc/symm.cu)
        "sdml_symm_alloc": [L, P],
        "sdml_symm_free": [P],
        "sdml_symm_export": [P, P],
        "sdml_symm_import": [P, P],
        "sdml_symm_unimport": [P],
        # dst, dpitch, src, spitch, width, height, stream
        "sdml_memcpy2d_async": [P, L, P, L, L, L, P],
        # stage boundary (csrc/boundary.cu)
        # flag, cnt, add, timeout_ns, err_dev, err_host, code, stream
        "sdml_bnd_wait": [P, P, ctypes.c_uint, ctypes.c_ulonglong, P, P, ctypes.c_uint, P],
        # flag_peer, cnt, add, bump, stream
        "sdml_bnd_signal": [P, P, ctypes.c_uint, ctypes.c_uint, P],
        # dst_peer, src, bytes, flag_peer, cnt, add, ack, timeout_ns, err_dev, err_host, code, stream
        "sdml_bnd_push": [P, P, L, P, P, ctypes.c_uint, P, ctypes.c_ulonglong, P, P, ctypes.c_uint, P],
        # n_tensors, params**, grads**, moms**, sizes*, rank, world, peer_stage**, peer_flags**,
        # local_counter, epoch, lr, momentum, weight_decay, grad_scale, stream
        "sdml_dp_allreduce_sgd": [I, P, P, P, P, I, I, P, P, P, ctypes.c_uint, F, F, F, F, P],
        # n_tensors, params**, grads**, moms**, sizes*, lr, momentum, weight_decay, grad_scale, stream
        "sdml_dp_allreduce_sgd_static": [I, P, P, P, P, F, F, F, F, P],
        # A, B, C, M, N, K, lda, ldb, ldc, b_mn, bias, aux, ld_aux, relu, out_fp32, alpha,
        # ag_world, ag_rank, ag_src**, ag_flags**, ag_count, ag_max_chunks, ag_chunk_rows, epoch,
        # rs_world, rs_rank, rs_dst**, rs_ld, rs_done**, stream
        "sdml_gemm_bf16_comm": [P, P, P, I, I, I, L, L, L, I, P, P, L, I, I, F,
                                I, I, P, P, P, I, I, ctypes.c_uint, I, I, P, L, P, P],
        # landing, slots, slot_stride, ld_land, rows, cols, bias, aux, ld_aux, relu, out, ld_out, done, expected, stream
        "sdml_rs_finalize": [P, I, L, L, I, I, P, P, L, I, P, L, P, ctypes.c_uint, P],
    }
    extra = {
        # dz, x, M, N, K, ld_dz, ld_x, alpha, rs_world, rs_rank, rs_dst**, rs_ld, rs_done**, stream
        "sdml_wgrad_rs": [P, P, I, I, I, L, L, F, I, I, P, L, P, P],
        # landing, slots, slot_stride, ld_land, rows, cols, master, mom, ld_w, lr, momentum, wd,
        # wout**, ld_wout, world, done, expected, ready**, exit_counter, stream
        "sdml_rs_finalize_sgd": [P, I, L, L, I, I, P, P, L, F, F, F, P, L, I, P, ctypes.c_uint, P, P, P],
        # counter, expected, stream
        "sdml_wait_counter": [P, ctypes.c_uint, P],
        # n_tensors, params**, grads**, moms**, sizes*, rank, world, stage_odd**, stage_even**, flags**,
        # local_counter, epoch_dev, lr, momentum, weight_decay, grad_scale, stream
        "sdml_dp_allreduce_sgd_dev_epoch": [I, P, P, P, P, I, I, P, P, P, P, P, F, F, F, F, P],
        # n_tensors, params**, grads**, moms**, sizes*, rank, world, peer_buf**, epoch_dev, exit_counter,
        # lr, momentum, weight_decay, grad_scale, timeout_ns, err, stamps, stream
        "sdml_dp_ll_allreduce_sgd": [I, P, P, P, P, I, I, P, P, P, F, F, F, F, ctypes.c_ulonglong, P, P, P],
    }
    for name, argtypes in extra.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.restype = c_int
            fn.argtypes = argtypes
    lenet = {
        # x, w1, b1, w2, b2, B, z3, z1_save, idx1, idx2, mask2, drop_p, seed, offset_dev, offset_imm,
        # zero_this, bump_counter, bump_by, stream
        "sdml_lenet_s0_fwd": [P, P, P, P, P, I, P, P, P, P, P, F, L, P, L, P, P, L, P],
        # z3, targets, w1, b1, w2, b2, B, train, grad_scale, drop_p, seed, offset_dev, offset_imm,
        # loss_sum, loss_scale, correct, dz3, dz4_scratch, dw1, db1, dw2, db2, bump_counter, bump_by,
        # stream, wgrad_stream
        "sdml_lenet_s1": [P, P, P, P, P, P, I, I, F, F, L, P, L, P, F, P, P, P, P, P, P, P, P, L, P, P],
        # x, w2, z3, dz3, z1_save, idx1, idx2, mask2, B, dw1, db1, dw2, db2, stream
        "sdml_lenet_s0_bwd": [P, P, P, P, P, P, P, P, I, P, P, P, P, P],
    }
    for name, argtypes in lenet.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.restype = c_int
            fn.argtypes = argtypes
    fp8 = {
        # src, ld_src, rows, cols, dst, ld_dst, sf, stream
        "sdml_quant_mxfp8": [P, L, I, I, P, L, P, P],
        # A, sfa, B, sfb, C, M, N, K, lda, ldb, ldc, bias, aux, ld_aux, relu, out_fp32, accumulate, alpha, stream
        "sdml_gemm_mxfp8": [P, P, P, P, P, I, I, I, L, L, L, P, P, L, I, I, I, F, P],
        # ... same + bn before stream
        "sdml_gemm_mxfp8_2cta": [P, P, P, P, P, I, I, I, L, L, L, P, P, L, I, I, I, F, I, P],
        # src, ld_src, rows, cols, dst_row, ld_row, sf_row, dst_col, ld_col, sf_col, groups_col, stream
        "sdml_quant_mxfp8_dual": [P, L, I, I, P, L, P, P, L, P, I, P],
        # src, ld_src, rows, cols, dst_row, ld_row, sf_row, groups_row, dst_col, ld_col, sf_col, groups_col, stream
        "sdml_quant_mxfp8_dual_at": [P, L, I, I, P, L, P, I, P, L, P, I, P],
        # A, sfa, B, sfb, C, M, N, K, lda, ldb, ldc, bias, aux, ld_aux, relu, alpha,
        # q_row, ld_qrow, sf_qrow, q_col, ld_qcol, sf_qcol, groups_qcol, stream
        # A, sfa, B, sfb, M, N, K, lda, ldb, alpha, rs_world, rs_rank, rs_dst**, rs_ld, rs_done**, stream
        "sdml_gemm_mxfp8_2cta_rs": [P, P, P, P, I, I, I, L, L, F, I, I, P, L, P, P],
        "sdml_gemm_mxfp8_2cta_qout": [P, P, P, P, P, I, I, I, L, L, L, P, P, L, I, F, P, L, P, P, L, P, I, P],
    }
    for name, argtypes in fp8.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.restype = c_int
            fn.argtypes = argtypes
    p2p = {
        "sdml_p2p_nccl_version": [],
        "sdml_p2p_unique_id": [P],
        # uid128, nranks, rank, handle*
        "sdml_p2p_create": [P, I, I, P],
        # handle, abort
        "sdml_p2p_destroy": [P, I],
        # handle, ops*, n_ops, after_stream, order_after, ticket*
        "sdml_p2p_enqueue": [P, P, I, P, I, P],
        # handle, ticket, stream
        "sdml_p2p_wait_stream": [P, L, P],
        # handle, ticket, timeout_ms
        "sdml_p2p_wait_host": [P, L, I],
    }
    for name, argtypes in p2p.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.restype = c_int
            fn.argtypes = argtypes
    if hasattr(lib, "sdml_mxfp8_sf_bytes"):
        lib.sdml_mxfp8_sf_bytes.restype = c_longlong
        lib.sdml_mxfp8_sf_bytes.argtypes = [I, I]
    if hasattr(lib, "sdml_gemm_comm_grid"):
        lib.sdml_gemm_comm_grid.restype = c_int
        lib.sdml_gemm_comm_grid.argtypes = []
    for name, argtypes in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.restype = c_int
        fn.argtypes = argtypes


def cuda_lib(build: bool = True) -> ctypes.CDLL:
    """Return the loaded ``libsdml_b200.so`` (building it first when stale and nvcc exists)."""
    global _cuda_lib
    with _lock:
        if _cuda_lib is None:
            path = _build.CUDA_LIB
            if build:
                try:
                    _build.build_cuda()
                except Exception as e:
                    # a stale library would run old kernels against new Python / ctypes signatures
                    if not path.exists() or _build.cuda_sources_newer():
                        raise NativeError(f"rebuilding {path.name} failed and the sources are newer than the "
                                          f"library on disk: {e}") from e
            if not path.exists():
                raise NativeError(f"{path} is missing: run `python -m simple_distributed_machine_learning_b200.ops._build`")
            lib = ctypes.CDLL(str(path))
            _declare_cuda(lib)
            _cuda_lib = lib
        return _cuda_lib


def sched_lib() -> ctypes.CDLL:
    global _sched_lib
    with _lock:
        if _sched_lib is None:
            path = _build.build_sched()
            _sched_lib = ctypes.CDLL(str(path))
        return _sched_lib


_wait_err = None     # pinned {code, seen, wanted, 0} written by a kernel whose wait for a peer expired

_WAIT_CODES = {11: "all-gather flag of a tensor-parallel peer", 12: "reduce-scatter landing slots (tensor-parallel peers)",
               13: "weight-gradient landing slots (data-parallel replicas)",
               14: "refreshed-weights counter (data-parallel row owners)", 15: "staged DP all-reduce flags"}


def set_wait_timeout(timeout_s: float) -> None:
    """Bound the in-kernel waits of the fused collectives (process-wide, see sdml_set_wait_timeout)."""
    global _wait_err
    import torch
    if not torch.cuda.is_available():
        return
    if _wait_err is None:
        _wait_err = torch.zeros(4, dtype=torch.int32).pin_memory()
    cuda_lib().sdml_set_wait_timeout(int(max(0.0, timeout_s) * 1e9), _wait_err.data_ptr())


def check_waits() -> None:
    """Raise if a fused-collective kernel gave up waiting for another GPU (free: pinned host read)."""
    if _wait_err is not None and int(_wait_err[0]) != 0:
        code = int(_wait_err[0])
        raise NativeError(f"in-kernel wait timed out: {_WAIT_CODES.get(code, code)} "
                          f"(saw {int(_wait_err[1]) & 0xFFFFFFFF}, wanted {int(_wait_err[2]) & 0xFFFFFFFF}): "
                          "a peer GPU is dead or stalled")


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = cuda_lib().sdml_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed (rc={rc}): {msg}")


def last_error() -> str:
    return cuda_lib().sdml_last_error().decode(errors="replace")


def launch_count() -> int:
    return int(cuda_lib().sdml_launch_count())


def loaded_paths() -> list[Path]:
    out = []
    if _cuda_lib is not None:
        out.append(_build.CUDA_LIB)
    if _sched_lib is not None:
        out.append(_build.SCHED_LIB)
    return out
