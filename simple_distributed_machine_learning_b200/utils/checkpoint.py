"""Per-stage checkpoints.

The reference has no checkpoint code at all; the format honoured here is what
``torch.save(stage.state_dict())`` of the reference's own classes would produce, one
file per pipeline stage:

  stage0.pt : conv1.weight [10,1,5,5], conv1.bias [10], conv2.weight [20,10,5,5], conv2.bias [20]
  stage1.pt : fc1.weight [50,320], fc1.bias [50], fc2.weight [10,50], fc2.bias [10]

so ``Network1().load_state_dict(torch.load("stage0.pt"), strict=False)`` /
``Network2().load_state_dict(torch.load("stage1.pt"))`` work with the reference classes
(/root/reference/simple_distributed.py:26-31, 60-64; ``strict=False`` only because
``Network1`` wants an RPC peer for ``rref1``, which is not a parameter).
Beside each: ``stageN.optim.pt`` (momentum / Adam moments) and one ``meta.json``
(epoch, step, config).  TP shards add ``_tpK``; only DP replica 0 writes.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import torch


def stage_file(ckpt_dir: str, stage_idx: int, tp_idx: int = 0, tp: int = 1, suffix: str = "") -> str:
    name = f"stage{stage_idx}" + (f"_tp{tp_idx}" if tp > 1 else "") + suffix + ".pt"
    return os.path.join(ckpt_dir, name)


def save_stage(ckpt_dir: str, stage, pp_idx: int, optimizer=None, tp_idx: int = 0, tp: int = 1,
               write: bool = True) -> None:
    """Collect the stage's state and (``write``) store it.  EVERY data-parallel replica must call
    this -- a stage whose master weights / momentum are sharded over the replicas gathers them
    with a collective inside ``checkpoint_parts`` / ``optimizer_state`` -- but only replica 0
    passes ``write=True``."""
    parts = stage.checkpoint_parts(pp_idx)
    opt_sd = None
    if optimizer is not None:
        opt_sd = stage.optimizer_state() if hasattr(stage, "optimizer_state") else optimizer.state_dict()
    if not write:
        return
    os.makedirs(ckpt_dir, exist_ok=True)
    for idx, sd in parts.items():
        tmp = stage_file(ckpt_dir, idx, tp_idx, tp) + ".tmp"
        torch.save(sd, tmp)
        os.replace(tmp, stage_file(ckpt_dir, idx, tp_idx, tp))
    if opt_sd is not None:
        tmp = stage_file(ckpt_dir, pp_idx, tp_idx, tp, ".optim") + ".tmp"
        torch.save(opt_sd, tmp)
        os.replace(tmp, stage_file(ckpt_dir, pp_idx, tp_idx, tp, ".optim"))


def load_stage(ckpt_dir: str, stage, pp_idx: int, optimizer=None, tp_idx: int = 0, tp: int = 1) -> None:
    parts = {}
    for idx in stage.checkpoint_part_indices(pp_idx):
        parts[idx] = torch.load(stage_file(ckpt_dir, idx, tp_idx, tp), map_location="cpu", weights_only=True)
    stage.load_checkpoint_parts(pp_idx, parts)
    opt_path = stage_file(ckpt_dir, pp_idx, tp_idx, tp, ".optim")
    if optimizer is not None and os.path.exists(opt_path):
        optimizer.load_state_dict(torch.load(opt_path, map_location="cpu", weights_only=False))


def save_meta(ckpt_dir: str, epoch: int, step: int, config: Dict) -> None:
    os.makedirs(ckpt_dir, exist_ok=True)
    tmp = os.path.join(ckpt_dir, "meta.json.tmp")
    with open(tmp, "w") as f:
        json.dump({"epoch": epoch, "step": step, "config": config}, f, indent=1, default=str)
    os.replace(tmp, os.path.join(ckpt_dir, "meta.json"))


def load_meta(ckpt_dir: str) -> Optional[Dict]:
    path = os.path.join(ckpt_dir, "meta.json")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        return json.load(f)
