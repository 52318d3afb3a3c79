"""Config 1 of BASELINE.json on CPUs: the 2-stage reference model over gloo, world_size 2."""
import io, os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.multiprocessing as mp


def worker(rank, port, mb, sched, steps):
    torch.set_num_threads(1)
    from simple_distributed_machine_learning_b200.train import Trainer
    from simple_distributed_machine_learning_b200.utils.config import TrainConfig
    cfg = TrainConfig(rank=rank, world_size=2, interface="lo", master_addr="127.0.0.1", master_port=str(port),
                      device="cpu", backend="gloo", microbatches=mb, schedule=sched, epochs=1)
    t = Trainer(cfg, out=io.StringIO())
    it = list(t.train_loader)
    for i in range(10):
        t.train_batch(*it[i])
    t.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        loss = t.train_batch(*it[(10 + i) % len(it)])
        if loss is not None:
            float(loss)
    t.barrier()
    dt = time.perf_counter() - t0
    if rank == 1:
        print(f"ours cpu/gloo pp2 {sched} m={mb}: {60 * steps / dt:.0f} samples/s ({1e3 * dt / steps:.2f} ms/step)", flush=True)
    t.close()


if __name__ == "__main__":
    for mb, sched in ((1, "lockstep"), (4, "1f1b")):
        ctx = mp.get_context("spawn")
        ps = [ctx.Process(target=worker, args=(r, 29700 + mb, mb, sched, 80)) for r in range(2)]
        [p.start() for p in ps]
        [p.join() for p in ps]
