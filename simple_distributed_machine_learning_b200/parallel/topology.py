"""Rank <-> (pipeline, data, tensor) coordinates and process groups.

The reference hard-codes two ranks, ``"master"`` (rank 0) and ``"worker1"`` (rank 1)
(/root/reference/simple_distributed.py:34, 171, 179) and lets any further rank idle
(ref:177-184).  Here every rank has a role.  Rank order is tp-fastest, then dp, then
pp, so that the TP ranks of one stage are adjacent GPUs and a pipeline neighbour is
``dp*tp`` ranks away (all GPUs are one NVSwitch hop apart, so placement only matters
for readability, not bandwidth).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple


@dataclass(frozen=True)
class Topology:
    world_size: int
    pp: int
    dp: int = 1
    tp: int = 1

    def __post_init__(self):
        if self.pp * self.dp * self.tp != self.world_size:
            raise ValueError(
                f"pp({self.pp}) * dp({self.dp}) * tp({self.tp}) != world_size({self.world_size})")

    # ---- coordinates -------------------------------------------------------------------
    def coords(self, rank: int) -> Tuple[int, int, int]:
        """rank -> (pp_idx, dp_idx, tp_idx)."""
        if not (0 <= rank < self.world_size):
            raise ValueError(f"rank {rank} outside [0, {self.world_size})")
        tp_idx = rank % self.tp
        dp_idx = (rank // self.tp) % self.dp
        pp_idx = rank // (self.tp * self.dp)
        return pp_idx, dp_idx, tp_idx

    def rank_of(self, pp_idx: int, dp_idx: int = 0, tp_idx: int = 0) -> int:
        return (pp_idx * self.dp + dp_idx) * self.tp + tp_idx

    # ---- neighbours ----------------------------------------------------------------------
    def prev_stage_rank(self, rank: int) -> Optional[int]:
        s, d, t = self.coords(rank)
        return None if s == 0 else self.rank_of(s - 1, d, t)

    def next_stage_rank(self, rank: int) -> Optional[int]:
        s, d, t = self.coords(rank)
        return None if s == self.pp - 1 else self.rank_of(s + 1, d, t)

    def is_first_stage(self, rank: int) -> bool:
        return self.coords(rank)[0] == 0

    def is_last_stage(self, rank: int) -> bool:
        return self.coords(rank)[0] == self.pp - 1

    # ---- groups --------------------------------------------------------------------------
    def dp_group_ranks(self, rank: int) -> List[int]:
        s, _, t = self.coords(rank)
        return [self.rank_of(s, d, t) for d in range(self.dp)]

    def tp_group_ranks(self, rank: int) -> List[int]:
        s, d, _ = self.coords(rank)
        return [self.rank_of(s, d, t) for t in range(self.tp)]

    def pp_group_ranks(self, rank: int) -> List[int]:
        _, d, t = self.coords(rank)
        return [self.rank_of(s, d, t) for s in range(self.pp)]

    def all_dp_groups(self) -> List[List[int]]:
        return [[self.rank_of(s, d, t) for d in range(self.dp)]
                for s in range(self.pp) for t in range(self.tp)]

    def all_tp_groups(self) -> List[List[int]]:
        return [[self.rank_of(s, d, t) for t in range(self.tp)]
                for s in range(self.pp) for d in range(self.dp)]

    def all_pp_groups(self) -> List[List[int]]:
        return [[self.rank_of(s, d, t) for s in range(self.pp)]
                for d in range(self.dp) for t in range(self.tp)]

    def worker_name(self, rank: int) -> str:
        """The reference's RPC names (ref:171, 179), kept for log messages."""
        return "master" if rank == 0 else f"worker{rank}"

    def describe(self) -> str:
        parts = [f"pp{self.pp}"]
        if self.dp > 1:
            parts.append(f"dp{self.dp}")
        if self.tp > 1:
            parts.append(f"tp{self.tp}")
        return "x".join(parts)


@dataclass
class ProcessGroups:
    """torch.distributed groups for one rank (``None`` when the dimension has size 1)."""
    topo: Topology
    rank: int
    # boundary b joins pipeline stages b and b+1 of this rank's (dp, tp) column; one
    # 2-rank group per boundary AND per direction
    fwd_p2p: Dict[int, object] = None   # activations   stage b -> b+1
    bwd_p2p: Dict[int, object] = None   # gradients     stage b+1 -> b
    dp: object = None
    tp: object = None
    ctrl: object = None      # CPU (gloo) group for scalars / barriers / handle exchange


def build_process_groups(topo: Topology, rank: int, backend: str) -> ProcessGroups:
    """Create the sub-groups.  Every rank must call this with identical arguments.

    Each stage boundary gets its own pair of 2-rank communicators, one per direction.
    A rank therefore receives and sends on *different* communicators (hence different
    NCCL streams): a receive posted early for the next micro-batch can never sit in
    front of a send the downstream rank is waiting for, and activations never queue
    behind gradients (the classic bidirectional p2p dead-lock).
    """
    import torch.distributed as dist

    groups = ProcessGroups(topo=topo, rank=rank, fwd_p2p={}, bwd_p2p={})
    world = list(range(topo.world_size))
    _, my_d, my_t = topo.coords(rank)
    for d in range(topo.dp):
        for t in range(topo.tp):
            for b in range(topo.pp - 1):
                pair = [topo.rank_of(b, d, t), topo.rank_of(b + 1, d, t)]
                gf = dist.new_group(pair, backend=backend)
                gb = dist.new_group(pair, backend=backend)
                if d == my_d and t == my_t and rank in pair:
                    groups.fwd_p2p[b] = gf
                    groups.bwd_p2p[b] = gb
    if topo.dp > 1:
        for ranks in topo.all_dp_groups():
            g = dist.new_group(ranks, backend=backend)
            if rank in ranks:
                groups.dp = g
    if topo.tp > 1:
        for ranks in topo.all_tp_groups():
            g = dist.new_group(ranks, backend=backend)
            if rank in ranks:
                groups.tp = g
    if backend != "gloo":
        groups.ctrl = dist.new_group(world, backend="gloo")
    else:
        groups.ctrl = dist.group.WORLD
    return groups
