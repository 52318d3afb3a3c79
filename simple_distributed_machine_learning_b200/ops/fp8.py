"""MX-fp8 (E4M3 + UE8M0 scale per 32 elements) quantisation and block-scaled GEMM front-end
(``csrc/gemm_mxfp8.cu``: ``tcgen05.mma.kind::mxf8f6f4.block_scale``, scales in TMEM)."""
from __future__ import annotations

from dataclasses import dataclass

import torch

from . import _lib


@dataclass
class MXFP8Tensor:
    """Row-major fp8 payload + scale atoms in the layout the GEMM's ``tcgen05.cp`` expects."""
    data: torch.Tensor      # uint8 [rows, cols] (E4M3 bit patterns)
    scales: torch.Tensor    # uint8 [(cols/128) * ceil(rows/128) * 512]
    rows: int
    cols: int

    def dequantize(self) -> torch.Tensor:
        """fp32 [rows, cols] -- the oracle used by the tests."""
        vals = self.data.view(torch.float8_e4m3fn).float()
        r = torch.arange(self.rows, device=self.data.device)[:, None]
        b = torch.arange(self.cols // 32, device=self.data.device)[None, :]
        groups = (self.rows + 127) // 128
        off = ((b // 4) * groups + (r // 128)) * 512 + (r % 32) * 16 + ((r % 128) // 32) * 4 + (b % 4)
        e = self.scales[off.reshape(-1)].reshape(self.rows, self.cols // 32).to(torch.float32) - 127.0
        return vals * torch.exp2(e).repeat_interleave(32, dim=1)


def quantize_mxfp8(x: torch.Tensor) -> MXFP8Tensor:
    """bf16 [rows, cols] -> MX-fp8 along the contiguous (K) dimension; cols % 128 == 0."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    rows, cols = x.shape
    lib = _lib.cuda_lib()
    data = torch.empty(rows, cols, dtype=torch.uint8, device=x.device)
    nbytes = lib.sdml_mxfp8_sf_bytes(rows, cols)
    scales = torch.zeros(nbytes, dtype=torch.uint8, device=x.device)
    rc = lib.sdml_quant_mxfp8(x.data_ptr(), x.stride(0), rows, cols, data.data_ptr(), data.stride(0),
                              scales.data_ptr(), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "sdml_quant_mxfp8")
    return MXFP8Tensor(data, scales, rows, cols)


def gemm_mxfp8(A: MXFP8Tensor, B: MXFP8Tensor, C: torch.Tensor, *, bias=None, aux=None,
               relu: bool = False, alpha: float = 1.0) -> torch.Tensor:
    """C[M, N] = epilogue(alpha * deq(A)[M, K] . deq(B)[N, K]^T), fp32 accumulation."""
    M, K = A.rows, A.cols
    N = B.rows
    assert B.cols == K and C.shape == (M, N) and C.stride(1) == 1
    rc = _lib.cuda_lib().sdml_gemm_mxfp8(
        A.data.data_ptr(), A.scales.data_ptr(), B.data.data_ptr(), B.scales.data_ptr(), C.data_ptr(),
        M, N, K, A.data.stride(0), B.data.stride(0), C.stride(0),
        bias.data_ptr() if bias is not None else None, aux.data_ptr() if aux is not None else None,
        aux.stride(0) if aux is not None else 0, int(relu), int(C.dtype == torch.float32), float(alpha),
        torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "sdml_gemm_mxfp8")
    return C
