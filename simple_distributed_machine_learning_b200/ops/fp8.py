"""MX-fp8 (E4M3 + UE8M0 scale per 32 elements) quantisation and block-scaled GEMM front-end
(``csrc/gemm_mxfp8.cu``: ``tcgen05.mma.kind::mxf8f6f4.block_scale``, scales in TMEM).

The fp8 form of the reference's Linear+bias+ReLU layers and of their autograd backward
(/root/reference/simple_distributed.py:63-64, :75-77; fp32 on the CPU there)."""
from __future__ import annotations

import os
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


# ---------------------------------------------------------------------------------------------
# CPU emulation: the same arithmetic and the same byte layouts as csrc/gemm_mxfp8.cu, in plain
# PyTorch.  It lets the fp8 training plumbing (dual copies, step-wide transposed buffers, K-slices,
# accumulate flags) run in the CPU test-suite, and it is the oracle the GPU quantisers are compared
# with.  Not a performance path.
# ---------------------------------------------------------------------------------------------
def _scale_offsets(rows: int, cols: int, device) -> torch.Tensor:
    """Byte offset of the scale of (row r, 32-column block b) inside the atom array."""
    r = torch.arange(rows, device=device)[:, None]
    b = torch.arange(cols // 32, device=device)[None, :]
    groups = (rows + 127) // 128
    return ((b // 4) * groups + (r // 128)) * 512 + (r % 32) * 16 + ((r % 128) // 32) * 4 + (b % 4)


def _emulate_quant(x: torch.Tensor, data_out: torch.Tensor, scales_out: torch.Tensor) -> None:
    """x [rows, cols] (any float dtype) -> E4M3 bytes into data_out [rows, cols] (may be a strided
    view) + UE8M0 exponents into scales_out (atom layout), blocks of 32 along the columns."""
    rows, cols = x.shape
    blk = x.float().reshape(rows, cols // 32, 32)
    amax = blk.abs().amax(dim=2)
    e = torch.where(amax > 0, torch.ceil(torch.log2(amax.clamp_min(1e-45) * (1.0 / 448.0))),
                    torch.full_like(amax, -127.0)).clamp_(-127, 127)
    q = (blk * torch.exp2(-e)[:, :, None]).clamp_(-448.0, 448.0).to(torch.float8_e4m3fn)
    data_out.copy_(q.reshape(rows, cols).view(torch.uint8))
    scales_out[_scale_offsets(rows, cols, x.device).reshape(-1)] = (e + 127).to(torch.uint8).reshape(-1)


def quantize_mxfp8(x: torch.Tensor, out: "MXFP8Tensor" = None) -> MXFP8Tensor:
    """bf16 [rows, cols] -> MX-fp8 along the contiguous (K) dimension; cols % 128 == 0.
    ``out``: re-quantise in place into an existing tensor of the same shape (weights after an
    optimizer step: consumers -- and CUDA graphs -- keep pointing at the same memory)."""
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    rows, cols = x.shape
    assert cols % 128 == 0
    if out is not None:
        assert (out.rows, out.cols) == (rows, cols)
        data, scales = out.data, out.scales
    else:
        data = torch.empty(rows, cols, dtype=torch.uint8, device=x.device)
        scales = torch.zeros(sf_bytes(rows, cols), dtype=torch.uint8, device=x.device)
    if not x.is_cuda:
        _emulate_quant(x, data, scales)
        return out if out is not None else MXFP8Tensor(data, scales, rows, cols)
    lib = _lib.cuda_lib()
    rc = lib.sdml_quant_mxfp8(x.data_ptr(), x.stride(0), rows, cols, data.data_ptr(), data.stride(0),
                              scales.data_ptr(), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "sdml_quant_mxfp8")
    return out if out is not None else MXFP8Tensor(data, scales, rows, cols)


def sf_bytes(rows: int, cols: int) -> int:
    return ((cols + 127) // 128) * ((rows + 127) // 128) * 512


class MXFP8Transposed:
    """Step-wide TRANSPOSED fp8 copy of a row-stacked [total_rows, cols] bf16 buffer: payload
    [cols, total_rows] scaled along the original rows, filled one row range (micro-batch) at a
    time by ``quantize_mxfp8_dual`` and consumed as the K-major operand of a GEMM that contracts
    over a row range (the weight gradient of a group of micro-batches)."""

    def __init__(self, total_rows: int, cols: int, device):
        assert total_rows % 128 == 0 and cols % 128 == 0
        self.total_rows, self.cols = total_rows, cols
        self.data = torch.empty(cols, total_rows, dtype=torch.uint8, device=device)
        self.groups = (cols + 127) // 128
        self.scales = torch.zeros(sf_bytes(cols, total_rows), dtype=torch.uint8, device=device)

    def rows_slice(self, lo: int, hi: int) -> MXFP8Tensor:
        """Operand [cols, hi-lo] (K = hi-lo) over original rows lo..hi (multiples of 128)."""
        assert lo % 128 == 0 and hi % 128 == 0
        k0 = (lo // 128) * self.groups * 512
        k1 = (hi // 128) * self.groups * 512
        return MXFP8Tensor(self.data[:, lo:hi], self.scales[k0:k1], self.cols, hi - lo)


def quantize_mxfp8_dual(x: torch.Tensor, *, want_row: bool = True, col_into: "MXFP8Transposed" = None,
                        row_offset: int = 0, want_col: bool = False, row_out: "MXFP8Tensor" = None,
                        col_out: "MXFP8Tensor" = None):
    """One read of bf16 ``x`` [rows, cols] -> (row-wise MXFP8Tensor or None, transposed operand or None).

    ``col_into`` + ``row_offset`` write the transposed copy into rows-range
    [row_offset, row_offset + rows) of a step-wide ``MXFP8Transposed``; ``want_col`` without
    ``col_into`` returns a stand-alone transposed ``MXFP8Tensor`` [cols, rows].  ``row_out`` /
    ``col_out``: overwrite existing tensors of the right shape instead of allocating."""
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    rows, cols = x.shape
    assert rows % 128 == 0 and cols % 128 == 0, "dual quantiser needs multiples of 128"
    if not x.is_cuda:
        return _emulate_dual(x, want_row, col_into, row_offset, want_col, row_out, col_out)
    lib = _lib.cuda_lib()
    row_t = None
    d_row = s_row = None
    ld_row = 0
    if row_out is not None:
        assert (row_out.rows, row_out.cols) == (rows, cols)
        row_t = row_out
        d_row, s_row, ld_row = row_t.data.data_ptr(), row_t.scales.data_ptr(), row_t.data.stride(0)
    elif want_row:
        data = torch.empty(rows, cols, dtype=torch.uint8, device=x.device)
        scales = torch.empty(sf_bytes(rows, cols), dtype=torch.uint8, device=x.device)
        row_t = MXFP8Tensor(data, scales, rows, cols)
        d_row, s_row, ld_row = data.data_ptr(), scales.data_ptr(), data.stride(0)
    col_t = None
    d_col = s_col = None
    ld_col = groups_col = 0
    if col_into is not None:
        assert col_into.cols == cols and row_offset % 128 == 0 and row_offset + rows <= col_into.total_rows
        d_col = col_into.data.data_ptr() + row_offset
        s_col = col_into.scales.data_ptr() + (row_offset // 128) * col_into.groups * 512
        ld_col, groups_col = col_into.data.stride(0), col_into.groups
    elif col_out is not None:
        assert (col_out.rows, col_out.cols) == (cols, rows)
        col_t = col_out
        d_col, s_col, ld_col, groups_col = (col_t.data.data_ptr(), col_t.scales.data_ptr(), col_t.data.stride(0),
                                            (cols + 127) // 128)
    elif want_col:
        data = torch.empty(cols, rows, dtype=torch.uint8, device=x.device)
        scales = torch.empty(sf_bytes(cols, rows), dtype=torch.uint8, device=x.device)
        col_t = MXFP8Tensor(data, scales, cols, rows)
        d_col, s_col, ld_col, groups_col = data.data_ptr(), scales.data_ptr(), data.stride(0), (cols + 127) // 128
    rc = lib.sdml_quant_mxfp8_dual(x.data_ptr(), x.stride(0), rows, cols, d_row, ld_row, s_row,
                                   d_col, ld_col, s_col, groups_col, torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "sdml_quant_mxfp8_dual")
    return row_t, col_t


def _emulate_dual(x, want_row, col_into, row_offset, want_col, row_out, col_out):
    rows, cols = x.shape
    row_t = None
    if row_out is not None or want_row:
        row_t = quantize_mxfp8(x, out=row_out)
    col_t = None
    xt = x.t().contiguous()                                   # [cols, rows]: blocks run along original rows
    if col_into is not None:
        assert col_into.cols == cols and row_offset % 128 == 0 and row_offset + rows <= col_into.total_rows
        k0 = (row_offset // 128) * col_into.groups * 512
        k1 = ((row_offset + rows) // 128) * col_into.groups * 512
        _emulate_quant(xt, col_into.data[:, row_offset:row_offset + rows], col_into.scales[k0:k1])
    elif col_out is not None or want_col:
        col_t = quantize_mxfp8(xt, out=col_out)
    return row_t, col_t


# tile: "1cta" = 128x128 single-CTA kernel, "2cta192" / "2cta128" = CTA-pair kernel, 256 x BN tile
_DEFAULT_TILE = os.environ.get("SDML_FP8_TILE", "2cta192")


# Fused output quantisation in the GEMM epilogue (sdml_gemm_mxfp8_2cta_qout).  Verified
# bit-identical to the quantiser pass on a B200 (tests/test_fp8_gpu.py), but MEASURED SLOWER: the
# 32x32 byte transposes of the epilogue outlast the main loop of the next tile (8-layer MLP step
# 26.2 ms with the separate dual-quantiser pass vs 42.8 ms fused, profiles/fp8_fused_quant_ab_r2.log).
# So the default stays the separate pass; SDML_FP8_FUSED_QUANT=1 selects the fused epilogue.
_FUSED_QOUT = os.environ.get("SDML_FP8_FUSED_QUANT", "0") == "1"


def gemm_mxfp8(A: MXFP8Tensor, B: MXFP8Tensor, C: torch.Tensor, *, bias=None, aux=None,
               relu: bool = False, alpha: float = 1.0, accumulate: bool = False,
               tile: str = None, q_row_out: MXFP8Tensor = None, q_col_into: "MXFP8Transposed" = None,
               q_row_offset: int = 0) -> torch.Tensor:
    """C[M, N] (+)= epilogue(alpha * deq(A)[M, K] . deq(B)[N, K]^T), fp32 accumulation.

    ``q_row_out`` / ``q_col_into`` (+ ``q_row_offset``): also produce the MX-fp8 copies of the
    bf16 result that the next GEMMs consume -- C quantised along N into ``q_row_out`` and C^T
    quantised along M into rows [q_row_offset, q_row_offset + M) of ``q_col_into`` -- from the
    GEMM's own epilogue when the fused kernel applies, else by a quantiser pass over C."""
    M, K = A.rows, A.cols
    N = B.rows
    assert B.cols == K and C.shape == (M, N) and C.stride(1) == 1
    want_q = q_row_out is not None or q_col_into is not None
    if want_q:
        assert C.dtype == torch.bfloat16 and not accumulate
        fused = (_FUSED_QOUT and C.is_cuda and M % 128 == 0 and N % 128 == 0 and M >= 256
                 and (tile or _DEFAULT_TILE).startswith("2cta"))
        if not fused:
            gemm_mxfp8(A, B, C, bias=bias, aux=aux, relu=relu, alpha=alpha, tile=tile)
            quantize_mxfp8_dual(C, want_row=False, row_out=q_row_out, col_into=q_col_into,
                                row_offset=q_row_offset)
            return C
        d_row = s_row = d_col = s_col = None
        ld_row = ld_col = groups_col = 0
        if q_row_out is not None:
            assert (q_row_out.rows, q_row_out.cols) == (M, N)
            d_row, s_row, ld_row = q_row_out.data.data_ptr(), q_row_out.scales.data_ptr(), q_row_out.data.stride(0)
        if q_col_into is not None:
            assert q_col_into.cols == N and q_row_offset % 128 == 0 and q_row_offset + M <= q_col_into.total_rows
            d_col = q_col_into.data.data_ptr() + q_row_offset
            s_col = q_col_into.scales.data_ptr() + (q_row_offset // 128) * q_col_into.groups * 512
            ld_col, groups_col = q_col_into.data.stride(0), q_col_into.groups
        rc = _lib.cuda_lib().sdml_gemm_mxfp8_2cta_qout(
            A.data.data_ptr(), A.scales.data_ptr(), B.data.data_ptr(), B.scales.data_ptr(), C.data_ptr(),
            M, N, K, A.data.stride(0), B.data.stride(0), C.stride(0),
            bias.data_ptr() if bias is not None else None, aux.data_ptr() if aux is not None else None,
            aux.stride(0) if aux is not None else 0, int(relu), float(alpha),
            d_row, ld_row, s_row, d_col, ld_col, s_col, groups_col, torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_gemm_mxfp8_2cta_qout")
        return C
    if not C.is_cuda:
        r = alpha * (A.dequantize() @ B.dequantize().t())
        if bias is not None:
            r = r + bias
        if relu:
            r = torch.relu(r)
        if aux is not None:
            r = r * (aux.float() > 0)
        if accumulate:
            C += r.to(C.dtype)
        else:
            C.copy_(r)
        return C
    tile = tile or _DEFAULT_TILE
    if tile.startswith("2cta") and M >= 256 and N >= 128:
        rc = _lib.cuda_lib().sdml_gemm_mxfp8_2cta(
            A.data.data_ptr(), A.scales.data_ptr(), B.data.data_ptr(), B.scales.data_ptr(), C.data_ptr(),
            M, N, K, A.data.stride(0), B.data.stride(0), C.stride(0),
            bias.data_ptr() if bias is not None else None, aux.data_ptr() if aux is not None else None,
            aux.stride(0) if aux is not None else 0, int(relu), int(C.dtype == torch.float32), int(accumulate),
            float(alpha), int(tile[4:] or 192), torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "sdml_gemm_mxfp8_2cta")
        return C
    rc = _lib.cuda_lib().sdml_gemm_mxfp8(
        A.data.data_ptr(), A.scales.data_ptr(), B.data.data_ptr(), B.scales.data_ptr(), C.data_ptr(),
        M, N, K, A.data.stride(0), B.data.stride(0), C.stride(0),
        bias.data_ptr() if bias is not None else None, aux.data_ptr() if aux is not None else None,
        aux.stride(0) if aux is not None else 0, int(relu), int(C.dtype == torch.float32), int(accumulate),
        float(alpha), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "sdml_gemm_mxfp8")
    return C
