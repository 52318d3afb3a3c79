"""Native NCCL point-to-point engine (csrc/p2p_engine.cu) on one GPU: a one-rank communicator
sending to itself exercises rendezvous, the side stream, producer ordering, tickets and the
bounded host wait.  The two-GPU pipeline run is in tests/test_multigpu.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_self_exchange_is_ordered_after_the_producer_and_ticketed():
    from simple_distributed_machine_learning_b200.ops import _lib
    from simple_distributed_machine_learning_b200.parallel.p2p_native import P2PChannel
    assert _lib.cuda_lib().sdml_p2p_nccl_version() >= 22000
    dev = torch.device("cuda", 0)
    ch = P2PChannel(P2PChannel.new_uid(), 1, 0, dev)
    try:
        src = torch.empty(1 << 20, dtype=torch.float32, device=dev)
        dst = torch.zeros_like(src)
        tickets = []
        for it in range(4):
            # the send must see the value written by the kernel queued just before it
            big = torch.randn(4096, 4096, device=dev)
            for _ in range(4):
                big = big @ big * 1e-3          # keeps the compute stream busy ahead of the fill
            src.fill_(float(it + 1))
            t = ch.exchange([(True, 0, src), (False, 0, dst)])
            tickets.append(t)
            ch.wait(t)                           # current stream waits; host does not
            got = dst.clone()
            assert ch.wait_host(t, 30.0)
            assert torch.all(got == float(it + 1))
        assert tickets == [0, 1, 2, 3]
        with pytest.raises(_lib.NativeError):
            ch.send(src, 3)                      # peer outside the communicator
    finally:
        ch.close()
