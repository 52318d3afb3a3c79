"""Python face of the native NCCL point-to-point engine (``csrc/p2p_engine.cu``).

One :class:`P2PChannel` = one NCCL communicator for ONE direction of ONE stage boundary with its
own side stream; transfers are ordered against the compute stream by CUDA events inside the
engine, so nothing here blocks the host.  Replaces the reference's blocking RPC hop
(/root/reference/simple_distributed.py:47-49, :80) and dist-autograd's gradient RPCs (ref:112).

The 128-byte NCCL rendezvous token of every channel is created by the sending side and spread
with one ``all_gather_object`` on the (gloo) control group -- the same out-of-band path the
symmetric-memory handles use.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib


class _Op(ctypes.Structure):
    _fields_ = [("is_send", ctypes.c_int), ("peer", ctypes.c_int), ("ptr", ctypes.c_void_p),
                ("bytes", ctypes.c_ulonglong)]


class P2PError(RuntimeError):
    pass


class P2PChannel:
    """A communicator of ``nranks`` (normally 2) ranks; this process is ``rank`` in it."""

    def __init__(self, uid: bytes, nranks: int, rank: int, device: torch.device):
        assert len(uid) == 128
        self.lib = _lib.cuda_lib()
        self.device = device
        self.nranks, self.rank = nranks, rank
        self._h = ctypes.c_void_p()
        with torch.cuda.device(device):
            rc = self.lib.sdml_p2p_create(uid, nranks, rank, ctypes.byref(self._h))
        _lib.check(rc, "sdml_p2p_create")

    @staticmethod
    def new_uid() -> bytes:
        buf = ctypes.create_string_buffer(128)
        _lib.check(_lib.cuda_lib().sdml_p2p_unique_id(buf), "sdml_p2p_unique_id")
        return buf.raw

    def _enqueue(self, ops: Sequence[Tuple[bool, int, torch.Tensor]], after_stream) -> int:
        arr = (_Op * len(ops))()
        for i, (is_send, peer, t) in enumerate(ops):
            assert t.is_cuda and t.is_contiguous()
            arr[i] = _Op(int(is_send), peer, t.data_ptr(), t.numel() * t.element_size())
        ticket = ctypes.c_longlong(-1)
        # (the default stream's handle is 0, so "no ordering" needs its own flag)
        rc = self.lib.sdml_p2p_enqueue(self._h, arr, len(ops), after_stream or None,
                                       int(after_stream is not None), ctypes.byref(ticket))
        _lib.check(rc, "sdml_p2p_enqueue")
        return ticket.value

    def send(self, t: torch.Tensor, peer: int, after: Optional[torch.cuda.Stream] = None) -> int:
        """Send once everything queued on ``after`` (default: the current stream) has run."""
        s = after if after is not None else torch.cuda.current_stream(self.device)
        return self._enqueue([(True, peer, t)], s.cuda_stream)

    def recv(self, t: torch.Tensor, peer: int, after: Optional[torch.cuda.Stream] = None,
             ordered: bool = True) -> int:
        """Receive into ``t``.  ``ordered`` (default): the receive kernel starts only after what is
        queued on ``after`` (default: the current stream) -- the buffer may still be read by
        earlier work (last step's weight-gradient GEMM reads the activation slot this lands in),
        and a receive kernel that starts early would sit on an SM, spinning, in the way of the
        persistent GEMM CTAs (measured: unordered receives cost the 2-stage MLP pipeline 20 %)."""
        if not ordered:
            return self._enqueue([(False, peer, t)], None)
        s = after if after is not None else torch.cuda.current_stream(self.device)
        return self._enqueue([(False, peer, t)], s.cuda_stream)

    def exchange(self, ops: Sequence[Tuple[bool, int, torch.Tensor]],
                 after: Optional[torch.cuda.Stream] = None) -> int:
        """Several sends / receives as one NCCL group (needed for a send+receive with oneself)."""
        s = after if after is not None else torch.cuda.current_stream(self.device)
        return self._enqueue(list(ops), s.cuda_stream)

    def wait(self, ticket: int, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Device-side wait: ``stream`` continues after the transfer; the host returns at once."""
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        _lib.check(self.lib.sdml_p2p_wait_stream(self._h, ticket, s.cuda_stream), "sdml_p2p_wait_stream")

    def wait_host(self, ticket: int, timeout_s: float) -> bool:
        """Bounded host wait; True when complete, False on time-out; raises on NCCL errors."""
        rc = self.lib.sdml_p2p_wait_host(self._h, ticket, int(timeout_s * 1000))
        if rc < 0:
            raise P2PError(_lib.last_error())
        return rc == 0

    def close(self, abort: bool = False) -> None:
        if self._h:
            self.lib.sdml_p2p_destroy(self._h, int(abort))
            self._h = ctypes.c_void_p()


class PipelineChannels:
    """The channels one pipeline rank needs: activations to the next stage / from the previous one,
    gradients the other way.  Built collectively by ALL ranks of the job (``ctrl_group``)."""

    def __init__(self, topo, rank: int, ctrl_group, device: torch.device, warmup_timeout_s: float = 120.0):
        self.prev_rank = topo.prev_stage_rank(rank)
        self.next_rank = topo.next_stage_rank(rank)
        # the sender of a channel creates its token: (src, dst) -> uid
        mine: Dict[Tuple[int, int], bytes] = {}
        if self.next_rank is not None:
            mine[(rank, self.next_rank)] = P2PChannel.new_uid()      # activations out
        if self.prev_rank is not None:
            mine[(rank, self.prev_rank)] = P2PChannel.new_uid()      # gradients out
        world = dist.get_world_size(group=ctrl_group)
        gathered: List[Optional[dict]] = [None] * world
        dist.all_gather_object(gathered, mine, group=ctrl_group)
        uids: Dict[Tuple[int, int], bytes] = {}
        for d in gathered:
            uids.update(d)
        # Create in one global order (sorted by (src, dst)) so that the blocking ncclCommInitRank
        # calls pair up instead of dead-locking: every channel involves exactly two ranks, and
        # both walk the same sorted list.
        self.act_out = self.act_in = self.grad_out = self.grad_in = None
        token = torch.zeros(8, dtype=torch.float32, device=device)
        for (src, dst) in sorted(uids):
            if rank not in (src, dst):
                continue
            ch = P2PChannel(uids[(src, dst)], 2, 0 if rank == src else 1, device)
            # NCCL connects two ranks lazily inside their first send/recv, and that set-up blocks
            # the host until the peer shows up: do it now, in the same global order, not in the
            # middle of a schedule where the two neighbours touch their channels in different orders
            t = ch.send(token, 1) if rank == src else ch.recv(token, 0, ordered=False)
            if not ch.wait_host(t, warmup_timeout_s):
                raise P2PError(f"rank {rank}: channel {src}->{dst} did not connect within {warmup_timeout_s:.0f}s")
            if (src, dst) == (rank, self.next_rank):
                self.act_out = ch
            elif (src, dst) == (rank, self.prev_rank):
                self.grad_out = ch
            elif (src, dst) == (self.prev_rank, rank):
                self.act_in = ch
            else:
                self.grad_in = ch

    def close(self, abort: bool = False) -> None:
        for ch in (self.act_out, self.act_in, self.grad_out, self.grad_in):
            if ch is not None:
                ch.close(abort)
