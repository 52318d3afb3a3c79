"""bf16 GEMM on tcgen05/TMEM/TMA (``csrc/gemm_tcgen05.cu``) -- Python front-end.

``gemm_bf16`` is a thin, allocation-free wrapper over the C entry point
``sdml_gemm_bf16``; the autograd-level Linear lives in :mod:`.linear`.  It is the GEMM under
the reference's ``fc1`` / ``fc2`` layers and their autograd backward
(/root/reference/simple_distributed.py:63-64, :75-77), which run on ATen's CPU ``addmm`` there.
"""
from __future__ import annotations

import torch

from . import _lib


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def gemm_bf16(A: torch.Tensor, B: torch.Tensor, C: torch.Tensor, *, a_mn: int = 0, b_mn: int = 0,
              bias: torch.Tensor | None = None, aux: torch.Tensor | None = None, relu: bool = False,
              accumulate: bool = False, alpha: float = 1.0, bn: int = 0) -> torch.Tensor:
    """C[M,N] = epilogue(alpha * op(A) . op(B)^T); see ``include/sdml_kernels.h``.

    ``a_mn=0``: A is [M,K]; ``a_mn=1``: A is [K,M] (e.g. dZ for wgrad).
    ``b_mn=0``: B is [N,K] (nn.Linear weight); ``b_mn=1``: B is [K,N].
    """
    assert A.is_cuda and B.is_cuda and C.is_cuda
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    assert A.dim() == 2 and B.dim() == 2 and C.dim() == 2
    assert A.stride(1) == 1 and B.stride(1) == 1 and C.stride(1) == 1
    M, N = C.shape
    K = A.shape[0] if a_mn else A.shape[1]
    assert (A.shape[1] if a_mn else A.shape[0]) == M, "A shape mismatch"
    assert (B.shape[0] if b_mn else B.shape[1]) == K and (B.shape[1] if b_mn else B.shape[0]) == N, "B shape mismatch"
    out_fp32 = C.dtype == torch.float32
    assert out_fp32 or C.dtype == torch.bfloat16
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N and bias.is_contiguous()
    if aux is not None:
        assert aux.dtype == torch.bfloat16 and aux.shape == C.shape and aux.stride(1) == 1
    lib = _lib.cuda_lib()
    rc = lib.sdml_gemm_bf16(
        A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, A.stride(0), B.stride(0), C.stride(0),
        int(a_mn), int(b_mn), bias.data_ptr() if bias is not None else None,
        aux.data_ptr() if aux is not None else None, aux.stride(0) if aux is not None else 0,
        int(relu), int(accumulate), int(out_fp32), float(alpha), int(bn), _stream_ptr())
    _lib.check(rc, "sdml_gemm_bf16")
    return C
