"""Worker functions for multi-process CPU (gloo) tests; must be importable by spawn."""
import io
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch  # noqa: E402


def _cfg(rank, world, port, **kw):
    from simple_distributed_machine_learning_b200.utils.config import TrainConfig
    base = dict(rank=rank, world_size=world, interface="lo", master_addr="127.0.0.1",
                master_port=str(port), device="cpu", backend="gloo", timeout_s=60.0)
    base.update(kw)
    return TrainConfig(**base)


def train_worker(rank, world, port, kw, out_dir, n_epochs_override=None):
    """Run Trainer.fit() and dump final stage weights + captured stdout."""
    torch.set_num_threads(1)
    from simple_distributed_machine_learning_b200.train import Trainer
    buf = io.StringIO()
    cfg = _cfg(rank, world, port, **kw)
    t = Trainer(cfg, out=buf)
    try:
        t.fit()
        sd = t.stage.checkpoint_parts(t.pp_idx)
        torch.save({"parts": sd, "log": buf.getvalue(), "coords": (t.pp_idx, t.dp_idx, t.tp_idx)},
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        t.close()


def dying_peer_worker(rank, world, port, kw, out_dir):
    """Rank 1 exits right after rendezvous; rank 0 must fail fast, not hang (ref:167 hangs)."""
    torch.set_num_threads(1)
    from simple_distributed_machine_learning_b200.train import Trainer
    from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineTimeout
    cfg = _cfg(rank, world, port, **kw)
    t = Trainer(cfg, out=io.StringIO())
    if rank == 1:
        os._exit(0)
    try:
        t.train_epoch(1)
        result = "completed"
    except PipelineTimeout as e:
        result = f"timeout: {e}"
    except Exception as e:  # gloo reports a closed connection
        result = f"error: {type(e).__name__}: {e}"
    with open(os.path.join(out_dir, "rank0.txt"), "w") as f:
        f.write(result)
    os._exit(0)


def gpu_train_worker(rank, world, port, kw, out_dir):
    """One rank per GPU (NCCL); dumps final weights like train_worker."""
    os.environ["LOCAL_RANK"] = str(rank)
    kw = dict(kw)
    os.environ.update(kw.pop("_env", {}))
    from simple_distributed_machine_learning_b200.train import Trainer
    from simple_distributed_machine_learning_b200.utils.config import TrainConfig
    buf = io.StringIO()
    base = dict(rank=rank, world_size=world, interface="lo", master_addr="127.0.0.1",
                master_port=str(port), device="cuda", backend="nccl", timeout_s=60.0)
    base.update(kw)
    t = Trainer(TrainConfig(**base), out=buf)
    try:
        t.fit()
        torch.cuda.synchronize()
        sd = t.stage.checkpoint_parts(t.pp_idx)
        torch.save({"parts": sd, "log": buf.getvalue(), "coords": (t.pp_idx, t.dp_idx, t.tp_idx),
                    "fused_dp": getattr(t.stage, "_fused_dp", None) is not None,
                    "native_p2p": getattr(t.engine, "native", None) is not None,
                    "p2p": t.engine.p2p_mode, "graph": bool(t.graph_ok), "graphs": len(t._graphs),
                    "direct_writes": getattr(t.engine.bnd, "direct_writes", 0) if t.engine.bnd is not None else 0},
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        t.close()


def stalled_peer_gpu_worker(rank, world, port, kw, out_dir):
    """Rank 1 stops issuing work after a few steps (a hung peer: it keeps its GPU memory mapped and
    simply sleeps); rank 0 must notice within --timeout_s through the in-kernel deadlines and raise
    PipelineTimeout instead of spinning forever (the reference disables every timeout, ref:36 / 167)."""
    import time
    os.environ["LOCAL_RANK"] = str(rank)
    from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineTimeout
    from simple_distributed_machine_learning_b200.train import Trainer
    from simple_distributed_machine_learning_b200.utils.config import TrainConfig
    base = dict(rank=rank, world_size=world, interface="lo", master_addr="127.0.0.1",
                master_port=str(port), device="cuda", backend="nccl")
    base.update(kw)
    t = Trainer(TrainConfig(**base), out=io.StringIO())
    t.stage.set_training(True)
    batches = list(t.train_loader)
    result, t0 = "completed", None
    try:
        for i in range(400):
            if rank == 1 and i == 6:
                torch.cuda.synchronize()
                open(os.path.join(out_dir, "rank1.stalled"), "w").write("x")
                time.sleep(600)                      # the parent test kills this process
            if rank == 0 and t0 is None and os.path.exists(os.path.join(out_dir, "rank1.stalled")):
                t0 = time.perf_counter()
            t.train_batch(*batches[i % len(batches)])
    except PipelineTimeout as e:
        result = f"timeout after {time.perf_counter() - (t0 or time.perf_counter()):.1f}s: {e}"
    except Exception as e:  # noqa: BLE001
        result = f"error: {type(e).__name__}: {e}"
    with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
        f.write(result)
    os._exit(0)
