"""The allocation-free entry point ``sdml_dp_allreduce_sgd_static``.

Who needs it: programs that embed this library but are not allowed to (or cannot afford
to) allocate GPU memory while running -- a host runtime that reserved the whole device
heap for its own arena, a step loop where ``cudaMalloc``'s device-wide synchronisation is
unacceptable, a deployment with a fixed memory plan.  For them the kernel's staging
buffers, flag and counter sit in a static array inside the library instead of in buffers
somebody has to ``cudaMalloc`` (single replica only: other GPUs cannot map a static array).

The test drives both entry points through the plain C ABI (ctypes, raw pointers) on the
same inputs and requires the same parameters / momentum / cleared gradients after each of
two consecutive calls -- the second call is the one that meets the array in the state the
first one left behind.  It keeps its own state (fresh tensors, library-internal epochs), so
it gives the same result alone in a fresh process and at the end of a full test run; the
file name makes pytest collect it last.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(10, 1, 5, 5), (10,), (20, 10, 5, 5), (20,), (50, 320), (50,), (10, 50), (10,)]   # the reference model


def _arrays(tensors):
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def _make(seed, device):
    g = torch.Generator(device=device).manual_seed(seed)
    params = [torch.randn(s, device=device, generator=g) for s in SHAPES]
    moms = [torch.randn(s, device=device, generator=g) * 0.1 for s in SHAPES]
    return params, moms


def test_static_scratch_matches_caller_provided_scratch_twice():
    from simple_distributed_machine_learning_b200.ops import _lib
    lib = _lib.cuda_lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    n = len(SHAPES)
    sizes = (ctypes.c_longlong * n)(*[int(torch.Size(s).numel()) for s in SHAPES])
    total = sum(sizes)
    lr, momentum, wd, gscale = 0.1, 0.5, 1e-3, 0.25

    p_ref, m_ref = _make(1, dev)
    p_new, m_new = [t.clone() for t in p_ref], [t.clone() for t in m_ref]
    # scratch for the existing entry point: provided (allocated) by the caller
    stage = [torch.zeros(total, device=dev), torch.zeros(total, device=dev)]
    flags = torch.zeros(8, dtype=torch.int32, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    flags_arr = (ctypes.c_void_p * 1)(flags.data_ptr())

    gen = torch.Generator(device=dev).manual_seed(7)
    for call in (1, 2):
        grads = [torch.randn(s, device=dev, generator=gen) for s in SHAPES]
        g_ref = [g.clone() for g in grads]
        g_new = [g.clone() for g in grads]
        stage_arr = (ctypes.c_void_p * 1)(stage[call & 1].data_ptr())
        rc = lib.sdml_dp_allreduce_sgd(n, _arrays(p_ref), _arrays(g_ref), _arrays(m_ref), sizes, 0, 1,
                                       stage_arr, flags_arr, ctypes.c_void_p(counter.data_ptr()),
                                       ctypes.c_uint(call), lr, momentum, wd, gscale, stream)
        _lib.check(rc, "sdml_dp_allreduce_sgd")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.sdml_dp_allreduce_sgd_static(n, _arrays(p_new), _arrays(g_new), _arrays(m_new), sizes,
                                              lr, momentum, wd, gscale, stream)
        _lib.check(rc, "sdml_dp_allreduce_sgd_static")
        torch.cuda.synchronize()
        if call == 2:   # set-up happened in call 1; from then on device memory must not move
            assert torch.cuda.mem_get_info()[0] == free_before
        for a, b in zip(p_new, p_ref):
            torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
        for a, b in zip(m_new, m_ref):
            torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
        for g in g_new:
            assert g.abs().max().item() == 0.0
        # and both agree with plain PyTorch SGD arithmetic on the same numbers
        if call == 1:
            p0, m0 = _make(1, dev)
            for p, m, g, pn in zip(p0, m0, grads, p_new):
                d = g * gscale + wd * p
                mv = momentum * m + d
                torch.testing.assert_close(pn, p - lr * mv, rtol=1e-5, atol=1e-6)


def test_static_scratch_rejects_oversized_calls():
    from simple_distributed_machine_learning_b200.ops import _lib
    lib = _lib.cuda_lib()
    dev = torch.device("cuda", 0)
    big = torch.zeros(65537, device=dev)
    arr = _arrays([big])
    sizes = (ctypes.c_longlong * 1)(big.numel())
    rc = lib.sdml_dp_allreduce_sgd_static(1, arr, arr, arr, sizes, 0.1, 0.5, 0.0, 1.0,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"65536" in lib.sdml_last_error()
    torch.cuda.synchronize()
    assert big.abs().max().item() == 0.0      # nothing was launched


@pytest.mark.skipif(torch.cuda.device_count() < 2 if torch.cuda.is_available() else True, reason="needs 2 GPUs")
def test_static_scratch_is_looked_up_per_device():
    """Every GPU that loads the library holds its own copy of the static array, and the entry point
    resolves the array's address per device on that device's first call.  Use it on GPU 0 and then on
    GPU 1 from the same process: each must update its own tensors correctly (a kernel handed GPU 0's
    address while running on GPU 1 would fault or scribble over the wrong memory)."""
    from simple_distributed_machine_learning_b200.ops import _lib
    lib = _lib.cuda_lib()
    n = len(SHAPES)
    sizes = (ctypes.c_longlong * n)(*[int(torch.Size(s).numel()) for s in SHAPES])
    lr, momentum, wd, gscale = 0.1, 0.5, 0.0, 1.0
    prev = torch.cuda.current_device()
    try:
        for ordinal in (0, 1, 0, 1):
            dev = torch.device("cuda", ordinal)
            torch.cuda.set_device(dev)
            params, moms = _make(11 + ordinal, dev)
            gen = torch.Generator(device=dev).manual_seed(3)
            grads = [torch.randn(s, device=dev, generator=gen) for s in SHAPES]
            want = [p - lr * (momentum * m + g * gscale) for p, m, g in zip(params, moms, grads)]
            rc = lib.sdml_dp_allreduce_sgd_static(n, _arrays(params), _arrays(grads), _arrays(moms), sizes,
                                                  lr, momentum, wd, gscale,
                                                  torch.cuda.current_stream(dev).cuda_stream)
            _lib.check(rc, "sdml_dp_allreduce_sgd_static")
            torch.cuda.synchronize(dev)
            for a, b in zip(params, want):
                torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
    finally:
        torch.cuda.set_device(prev)
