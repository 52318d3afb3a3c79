"""Adversarial workers for the hand-rolled cross-GPU flag protocols (spawned by test_stress_gpu.py).

compute-sanitizer's racecheck is not available on the GPU pool, so the protocols are stressed
instead: thousands of rounds, a random device-side delay on every rank before every protocol
step (the ranks drift apart by up to ~100 us in either direction), round counters that start just
below 2^32 so they wrap during the run, and payloads whose sum is known in closed form."""
import os
import random
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def _init(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ctrl = dist.new_group(list(range(world)), backend="gloo")
    return dev, ctrl


def _jitter(rng, max_us=100):
    # ~1.9 cycles per ns at 1.9 GHz; the exact figure does not matter
    torch.cuda._sleep(int(rng.random() * max_us * 1900))


class _FakeStage:
    act_dtype = torch.float32

    def __init__(self, width):
        self.width = width

    def in_shape(self, n):
        return (n, self.width)

    def out_shape(self, n):
        return (n, self.width)


def boundary_stress_worker(rank, world, port, kw, out_dir):
    """Rank 0 pushes activations (kernel copy <= 128 KiB, copy engine above), rank 1 checks them and
    pushes gradients back, `rounds` times, 4 micro-batches, counters starting at 2^32 - 5."""
    from simple_distributed_machine_learning_b200.parallel.boundary import SymmBoundary
    from simple_distributed_machine_learning_b200.parallel.topology import Topology
    dev, ctrl = _init(rank, world, port)
    rng = random.Random(1234 + rank)
    rounds, m, width = kw["rounds"], 4, kw["width"]
    rows = [kw["rows"], kw["rows"] // 2, kw["rows"], kw["rows"] // 2]
    topo = Topology(2, 2)
    bnd = SymmBoundary(topo, rank, ctrl, dev, _FakeStage(width), sum(rows), timeout_s=20.0)
    start = (1 << 32) - 5
    bnd.debug_set_round(start)
    dist.barrier(group=ctrl)
    got = torch.zeros(1, dtype=torch.float64, device=dev)
    bad = torch.zeros(1, dtype=torch.int64, device=dev)
    src = [torch.empty(r, width, device=dev) for r in rows]
    for step in range(rounds):
        bnd.begin_round(rows)
        for mb in range(m):
            val = float((step * m + mb) % 1021)
            if rank == 0:
                _jitter(rng)
                src[mb].fill_(val)
                bnd.send("act", mb, src[mb])
            else:
                _jitter(rng)
                bnd.wait("act", mb)
                x = bnd.in_slot("act", mb, (width,))
                bad += (x != val).sum()
                got += x.sum(dtype=torch.float64)
                src[mb].copy_(x).mul_(2.0)
                bnd.send("grad", mb, src[mb])
        if rank == 0:
            for mb in range(m):
                _jitter(rng, 30)
                bnd.wait("grad", mb)
                x = bnd.in_slot("grad", mb, (width,))
                val = 2.0 * float((step * m + mb) % 1021)
                bad += (x != val).sum()
                got += x.sum(dtype=torch.float64)
        bnd.end_round()
        if step % 64 == 63:
            torch.cuda.synchronize()
            bnd.check()
    torch.cuda.synchronize()
    bnd.check()
    expect = sum(float((s * m + mb) % 1021) * rows[mb] * width for s in range(rounds) for mb in range(m))
    expect *= 2.0 if rank == 0 else 1.0
    final_cnt = int(bnd.cnt[0].item()) & 0xFFFFFFFF
    torch.save({"bad": int(bad.item()), "got": float(got.item()), "expect": expect, "cnt": final_cnt,
                "start": start, "pushes": bnd.pushes}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier(group=ctrl)
    bnd.close()
    dist.destroy_process_group()


def dp_ll_stress_worker(rank, world, port, kw, out_dir):
    """`rounds` fused all-reduce + SGD steps (lr 1, no momentum) on gradients that are small integers,
    so the reduced update is exact in fp32 and the final parameters are known in closed form; epoch
    counter starts at 2^32 - 7 and wraps."""
    from simple_distributed_machine_learning_b200.parallel.symm import FusedDPSGD
    dev, ctrl = _init(rank, world, port)
    rng = random.Random(99 + rank)
    rounds = kw["rounds"]
    shapes = [(10, 1, 5, 5), (10,), (20, 10, 5, 5), (20,), (50, 320), (50,), (10, 50), (10,)]
    params = [torch.zeros(s, device=dev) for s in shapes]
    grads = [torch.zeros(s, device=dev) for s in shapes]
    moms = [torch.zeros(s, device=dev) for s in shapes]
    opt = FusedDPSGD(params, lambda: grads, moms, 1.0, 0.0, 0.0, list(range(world)), rank, ctrl, dev, timeout_s=20.0)
    opt.epoch_dev[0] = -7                  # uint32 2^32 - 7
    dist.barrier(group=ctrl)
    total = 0.0
    for step in range(rounds):
        _jitter(rng)
        for i, g in enumerate(grads):
            g.fill_(float(world * ((step + i + rank) % 5)))      # rank-dependent, mean over ranks is exact
        opt.step(grad_scale=1.0 / world)
        total += sum(float((step + r) % 5) for r in range(world))      # tensor 0's expected decrement
        if step % 128 == 127:
            torch.cuda.synchronize()
            opt.check()
    torch.cuda.synchronize()
    opt.check()
    exp = []
    for i in range(len(shapes)):
        exp.append(-sum(float((s + i + r) % 5) for s in range(rounds) for r in range(world)))
    ok = all(bool((p == e).all().item()) for p, e in zip(params, exp))
    zeroed = all(bool((g == 0).all().item()) for g in grads)
    torch.save({"ok": ok, "zeroed": zeroed, "p0": float(params[0].flatten()[0].item()), "e0": exp[0],
                "epoch": int(opt.epoch_dev[0].item()) & 0xFFFFFFFF,
                "flat": torch.cat([p.flatten() for p in params]).cpu()}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier(group=ctrl)
    dist.destroy_process_group()
