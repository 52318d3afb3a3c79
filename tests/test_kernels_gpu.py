"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference
of the same op.  Needs a B200: ``pytest -m gpu``."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from simple_distributed_machine_learning_b200.ops import _lib
    return _lib.cuda_lib()


def test_library_loaded_and_abi(lib):
    from simple_distributed_machine_learning_b200.ops import _lib
    assert lib.sdml_abi_version() == 1
    assert _lib.loaded_paths() and _lib.loaded_paths()[0].name == "libsdml_b200.so"


# ------------------------------------------------------------------ tcgen05 GEMM
def _gemm_oracle(A, B, a_mn, b_mn, bias, aux, relu, alpha, acc0):
    a = A.float().t() if a_mn else A.float()
    b = B.float().t() if b_mn else B.float()
    c = alpha * (a @ b.t())
    if bias is not None:
        c = c + bias
    if relu:
        c = torch.relu(c)
    if aux is not None:
        c = c * (aux.float() > 0)
    if acc0 is not None:
        c = c + acc0
    return c


GEMM_CASES = [
    # M, N, K, a_mn, b_mn, bias, aux, relu, fp32, acc, bn
    (128, 64, 64, 0, 0, False, False, False, False, False, 64),
    (128, 256, 256, 0, 0, False, False, False, False, False, 256),
    (256, 512, 320, 0, 1, False, False, False, False, False, 0),
    (256, 512, 320, 1, 0, False, False, False, False, False, 0),
    (256, 512, 320, 1, 1, False, False, False, True, False, 0),
    (1000, 776, 520, 0, 0, True, False, True, False, False, 0),     # ragged, forward epilogue
    (1000, 776, 520, 0, 1, False, True, False, False, False, 0),    # ragged, dgrad epilogue
    (776, 520, 1000, 1, 1, False, False, False, True, True, 0),     # ragged, wgrad accumulate
    (2048, 2048, 2048, 0, 0, True, False, True, False, False, 256),  # multi-wave persistent
    (60, 56, 320, 0, 0, True, False, True, False, False, 0),        # reference-sized fc1 (padded)
]


@pytest.mark.parametrize("M,N,K,a_mn,b_mn,bias,aux,relu,fp32,acc,bn", GEMM_CASES)
def test_gemm_bf16(M, N, K, a_mn, b_mn, bias, aux, relu, fp32, acc, bn):
    from simple_distributed_machine_learning_b200.ops.gemm import gemm_bf16
    g = torch.Generator(device="cuda").manual_seed(M + 3 * N + 7 * K)
    A = (torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g) * 0.5).bfloat16()
    B = (torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g) * 0.5).bfloat16()
    bias_t = torch.randn(N, device="cuda", generator=g) if bias else None
    aux_t = torch.randn(M, N, device="cuda", generator=g).bfloat16() if aux else None
    if acc:
        C = torch.randn(M, N, device="cuda", generator=g)
        acc0 = C.clone()
    else:
        C = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float32 if fp32 else torch.bfloat16)
        acc0 = None
    gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bias=bias_t, aux=aux_t, relu=relu, accumulate=acc,
              alpha=0.5 if acc else 1.0, bn=bn)
    ref = _gemm_oracle(A, B, a_mn, b_mn, bias_t, aux_t, relu, 0.5 if acc else 1.0, acc0)
    scale = max(ref.abs().max().item(), 1.0)
    tol = (2e-3 if fp32 else 1e-2) * scale
    assert torch.isfinite(C.float()).all()
    assert (C.float() - ref).abs().max().item() <= tol


BIG_GEMM_CASES = [
    # M, N, K, a_mn, b_mn, bn   (the training shapes of the 8-layer MLP and the full-size square)
    (2048, 8192, 8192, 0, 0, 0),       # forward of one micro-batch
    (2048, 8192, 8192, 0, 1, 0),       # dgrad
    (8192, 8192, 2048, 1, 1, 0),       # wgrad of one micro-batch (K = rows)
    (8192, 8192, 8192, 0, 0, 0),       # 128 k-blocks per tile: ring phase bits wrap many times
    (2048, 8192, 8192, 0, 0, 1128),    # CTA-pair 256 x 128 tile instantiation
    (1024, 4096, 8192, 0, 0, 256),     # single-CTA kernel on a deep K loop
]


@pytest.mark.parametrize("M,N,K,a_mn,b_mn,bn", BIG_GEMM_CASES)
def test_gemm_bf16_large_shapes_per_element(M, N, K, a_mn, b_mn, bn):
    """Full-size problems against an fp32 oracle (cuBLAS fp32 with TF32 off), bounded PER ELEMENT:
    |c - ref| <= 2 bf16 ulps of |ref| + the fp32-accumulation noise of a K-term sum."""
    from simple_distributed_machine_learning_b200.ops.gemm import gemm_bf16
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        g = torch.Generator(device="cuda").manual_seed(M + 3 * N + 7 * K + bn)
        A = (torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g) * 0.25).bfloat16()
        B = (torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g) * 0.25).bfloat16()
        fp32_out = bool(a_mn and b_mn)
        C = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float32 if fp32_out else torch.bfloat16)
        gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bn=bn)
        a = A.float().t() if a_mn else A.float()
        b = B.float() if b_mn else B.float().t()
        ref = a @ b
        err = (C.float() - ref).abs()
        # 2^-8 = one bf16 ulp (relative); accumulation noise ~ sqrt(K) * eps_fp32 * |a||b| per term
        bound = (0.0 if fp32_out else 2.0 * 2.0 ** -8) * ref.abs() + 4e-6 * math.sqrt(K) + 1e-5 * ref.abs()
        assert torch.isfinite(C.float()).all()
        bad = (err > bound).sum().item()
        assert bad == 0, f"{bad} of {M * N} elements outside the per-element bound; worst {float((err - bound).max()):.3e}"
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def test_gemm_rejects_bad_arguments():
    from simple_distributed_machine_learning_b200.ops.gemm import gemm_bf16
    from simple_distributed_machine_learning_b200.ops._lib import NativeError
    A = torch.zeros(128, 68, device="cuda", dtype=torch.bfloat16)[:, :60]   # lda = 68: not 16 B
    B = torch.zeros(64, 60, device="cuda", dtype=torch.bfloat16)
    C = torch.zeros(128, 64, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(NativeError):
        gemm_bf16(A, B, C)


# ------------------------------------------------------------------ softmax cross-entropy
@pytest.mark.parametrize("rows,classes,dtype", [(60, 10, torch.float32), (40, 10, torch.float32),
                                                (512, 8192, torch.bfloat16), (33, 1000, torch.bfloat16),
                                                (2048, 8192, torch.bfloat16),     # rows-in-registers path, 4 vectors / thread
                                                (700, 2048, torch.bfloat16),      # 1 vector / thread, more rows than CTAs
                                                (100, 4096, torch.float32),       # fp32 fast path
                                                (64, 16384, torch.bfloat16),      # too long for registers: generic path
                                                (37, 4104, torch.bfloat16)])      # 2 vectors / thread with a ragged tail
def test_softmax_ce(rows, classes, dtype):
    from simple_distributed_machine_learning_b200.ops.loss import softmax_ce
    g = torch.Generator(device="cuda").manual_seed(rows + classes)
    logits = (torch.randn(rows, classes, device="cuda", generator=g) * 3).to(dtype)
    t = torch.randint(0, classes, (rows,), device="cuda", generator=g)
    scale = 1.0 / rows
    loss_sum, dlogits, correct = softmax_ce(logits, t, grad_scale=scale, want_grad=True, want_correct=True)
    x = logits.float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(x, t, reduction="sum")
    (ref * scale).backward()
    assert loss_sum.item() == pytest.approx(ref.item(), rel=1e-4)
    tol = 1e-6 if dtype == torch.float32 else 2e-3 * scale * 4
    assert (dlogits.float() - x.grad).abs().max().item() <= max(tol, 1e-6)
    assert correct.item() == (logits.float().argmax(1) == t).sum().item()
    # accumulation across calls (one D2H read per eval epoch)
    softmax_ce(logits, t, want_grad=False, want_correct=True, loss_sum=loss_sum, correct=correct)
    assert loss_sum.item() == pytest.approx(2 * ref.item(), rel=1e-4)


def test_softmax_ce_strided_views_and_bad_targets():
    """Row-stacked buffers hand the kernel strided micro-batch views; a target outside [0, classes)
    must not index memory: the loss turns NaN, nothing else is touched."""
    from simple_distributed_machine_learning_b200.ops.loss import softmax_ce
    big = (torch.randn(256, 1024 + 64, device="cuda") * 2).bfloat16()
    logits = big[:, :1024]                                   # ld = 1088 elements (16-byte aligned rows)
    dl_big = torch.zeros_like(big)
    t = torch.randint(0, 1024, (256,), device="cuda")
    loss, dl, _ = softmax_ce(logits, t, grad_scale=1.0, dlogits=dl_big[:, :1024])
    x = logits.float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(x, t, reduction="sum")
    ref.backward()
    assert loss.item() == pytest.approx(ref.item(), rel=1e-4)
    assert (dl_big[:, :1024].float() - x.grad).abs().max().item() <= 8e-3
    assert dl_big[:, 1024:].abs().max().item() == 0.0        # nothing written past the view
    for classes in (1024, 10):                               # fast path and generic path
        lg = torch.randn(8, classes, device="cuda")
        bad = torch.randint(0, classes, (8,), device="cuda")
        bad[3] = classes + 5
        bad[5] = -1
        loss, dl, _ = softmax_ce(lg, bad, grad_scale=1.0)
        torch.cuda.synchronize()
        assert math.isnan(loss.item())
        p = torch.softmax(lg, 1)
        torch.testing.assert_close(dl[3], p[3], rtol=1e-4, atol=1e-6)    # no -1 term anywhere in that row


def test_cross_entropy_autograd_matches_torch():
    from simple_distributed_machine_learning_b200.ops.loss import cross_entropy
    logits = torch.randn(60, 10, device="cuda", requires_grad=True)
    t = torch.randint(0, 10, (60,), device="cuda")
    cross_entropy(logits, t, 1.0 / 60).backward()
    ref = logits.detach().clone().requires_grad_(True)
    torch.nn.functional.nll_loss(torch.log_softmax(ref, 1), t).backward()   # ref:79 + ref:111
    torch.testing.assert_close(logits.grad, ref.grad, rtol=1e-5, atol=1e-7)


# ------------------------------------------------------------------ optimizers
@pytest.mark.parametrize("momentum,wd", [(0.5, 0.0), (0.0, 0.0), (0.9, 1e-2)])
def test_fused_sgd_matches_torch(momentum, wd):
    from simple_distributed_machine_learning_b200.ops.optim import FusedSGD
    shapes = [(10, 1, 5, 5), (10,), (50, 320), (50,), (1037,), (3,)]
    g = torch.Generator(device="cuda").manual_seed(0)
    ps = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    ref_ps = [p.clone().requires_grad_(True) for p in ps]
    lowp = [torch.empty(s, device="cuda", dtype=torch.bfloat16) for s in shapes]
    grads = [torch.zeros_like(p) for p in ps]
    opt = FusedSGD(ps, lr=0.1, momentum=momentum, weight_decay=wd, grads=grads, lowp=lowp)
    ref = torch.optim.SGD(ref_ps, lr=0.1, momentum=momentum, weight_decay=wd)
    for step in range(4):
        for gbuf, rp in zip(grads, ref_ps):
            gr = torch.randn(gbuf.shape, device="cuda", generator=g)
            gbuf.copy_(gr * 4.0)            # un-scaled accumulator; grad_scale undoes the 4x
            rp.grad = gr.clone()
        opt.step(grad_scale=0.25, zero_grad=True)
        ref.step()
        for p, rp, lp, gbuf in zip(ps, ref_ps, lowp, grads):
            torch.testing.assert_close(p, rp.detach(), rtol=1e-5, atol=1e-6)
            assert torch.equal(lp, p.to(torch.bfloat16))
            assert gbuf.abs().max().item() == 0.0


def test_fused_adam_matches_torch():
    from simple_distributed_machine_learning_b200.ops.optim import FusedAdam
    shapes = [(64, 64), (64,), (1025,)]
    g = torch.Generator(device="cuda").manual_seed(1)
    ps = [torch.randn(s, device="cuda", generator=g) for s in shapes]
    ref_ps = [p.clone().requires_grad_(True) for p in ps]
    grads = [torch.zeros_like(p) for p in ps]
    opt = FusedAdam(ps, lr=1e-2, weight_decay=1e-2, grads=grads)
    ref = torch.optim.Adam(ref_ps, lr=1e-2, weight_decay=1e-2)
    for step in range(5):
        for gbuf, rp in zip(grads, ref_ps):
            gr = torch.randn(gbuf.shape, device="cuda", generator=g)
            gbuf.copy_(gr)
            rp.grad = gr.clone()
        opt.step()
        ref.step()
        for p, rp in zip(ps, ref_ps):
            torch.testing.assert_close(p, rp.detach(), rtol=2e-5, atol=2e-6)


# ------------------------------------------------------------------ small fp32 linear
@pytest.mark.parametrize("M,N,K,relu", [(60, 50, 320, True), (60, 10, 50, False), (40, 50, 320, True), (7, 10, 50, False)])
def test_linear_f32_matches_torch(M, N, K, relu):
    from simple_distributed_machine_learning_b200.ops.linear import linear_f32
    g = torch.Generator(device="cuda").manual_seed(M * N)
    x = torch.randn(M, K, device="cuda", generator=g, requires_grad=True)
    w = (torch.randn(N, K, device="cuda", generator=g) / math.sqrt(K)).requires_grad_(True)
    b = torch.randn(N, device="cuda", generator=g, requires_grad=True)
    y = linear_f32(x, w, b, relu=relu)
    dy = torch.randn(M, N, device="cuda", generator=g)
    y.backward(dy)
    xr, wr, br = (t.detach().clone().requires_grad_(True) for t in (x, w, b))
    yr = torch.nn.functional.linear(xr, wr, br)
    if relu:
        yr = torch.relu(yr)
    yr.backward(dy)
    torch.testing.assert_close(y, yr, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(w.grad, wr.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(b.grad, br.grad, rtol=1e-4, atol=1e-4)


def test_linear_f32_dropout_mask_is_consistent():
    from simple_distributed_machine_learning_b200.ops.linear import linear_f32
    x = torch.randn(60, 320, device="cuda", requires_grad=True)
    w = (torch.randn(50, 320, device="cuda") / 18).requires_grad_(True)
    b = torch.zeros(50, device="cuda", requires_grad=True)
    y = linear_f32(x, w, b, relu=True, drop_p=0.5, seed=7, offset=123)
    y0 = torch.relu(torch.nn.functional.linear(x, w, b)).detach()
    kept = (y != 0)
    active = y0 > 0
    frac = kept.sum().item() / max(1, active.sum().item())
    assert 0.4 < frac < 0.6                                  # p = 0.5 of the active units survive
    torch.testing.assert_close(y[kept], (2.0 * y0)[kept], rtol=1e-5, atol=1e-5)   # scaled by 1/(1-p)
    y.sum().backward()
    # backward regenerates the same mask: gradient w.r.t. bias counts kept units, scaled by 2
    torch.testing.assert_close(b.grad, 2.0 * kept.float().sum(0), rtol=1e-5, atol=1e-5)
    # a different offset gives a different mask
    off = torch.full((1,), 1 << 24, dtype=torch.int64, device="cuda")
    y2 = linear_f32(x, w, b, relu=True, drop_p=0.5, seed=7, offset=123, offset_dev=off)
    assert ((y2 != 0) != kept).any()


def test_colsum_and_cast(lib):
    from simple_distributed_machine_learning_b200.ops import _lib
    x = torch.randn(1000, 778, device="cuda").bfloat16()
    out = torch.full((778,), 5.0, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.sdml_colsum_bf16(x.data_ptr(), x.stride(0), 1000, 778, out.data_ptr(), 0, s), "colsum")
    torch.testing.assert_close(out, x.float().sum(0), rtol=1e-4, atol=1e-3)
    _lib.check(lib.sdml_colsum_bf16(x.data_ptr(), x.stride(0), 1000, 778, out.data_ptr(), 1, s), "colsum")
    torch.testing.assert_close(out, 2 * x.float().sum(0), rtol=1e-4, atol=2e-3)
    src = torch.randn(100003, device="cuda")
    dst = torch.empty(100003, device="cuda", dtype=torch.bfloat16)
    _lib.check(lib.sdml_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), s), "cast")
    assert torch.equal(dst, src.bfloat16())
