"""Latency of the fused DP all-reduce + SGD kernel alone (run under torchrun, >= 2 GPUs).

SDML_DP_STAMPS=1 adds the kernel's own globaltimer probes: per call, time from the first CTA's
entry to the last push issued and to the last CTA's exit (medians over the ring of 64 calls)."""
import os, sys, statistics
from pathlib import Path
import torch
import torch.distributed as dist
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.parallel.symm import FusedDPSGD

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
ctrl = dist.new_group(list(range(world)), backend="gloo")
shapes = [(10, 1, 5, 5), (10,), (20, 10, 5, 5), (20,), (50, 320), (50,), (10, 50), (10,)]
params = [torch.randn(s, device=dev) for s in shapes]
grads = [torch.randn(s, device=dev) for s in shapes]
moms = [torch.zeros(s, device=dev) for s in shapes]
opt = FusedDPSGD(params, lambda: grads, moms, 0.1, 0.5, 0.0, list(range(world)), rank, ctrl, dev, timeout_s=20.0)
for _ in range(20):
    opt.step(grad_scale=1.0 / world)
torch.cuda.synchronize(); dist.barrier()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(20):
        opt.step(grad_scale=1.0 / world)
for _ in range(3):
    g.replay()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    g.replay()
e1.record(); torch.cuda.synchronize()
opt.check()
line = f"rank {rank} world {world}: fused DP all-reduce+SGD {1e3 * e0.elapsed_time(e1) / 200:.2f} us per call"
if opt.stamps is not None:
    st = opt.stamps.view(64, 4).cpu()
    push = [int(r[1] - r[0]) for r in st if r[2] > 0 and r[0] > 0]
    tot = [int(r[2] - r[0]) for r in st if r[2] > 0 and r[0] > 0]
    if tot:
        line += (f"; in-kernel: enter->pushed {statistics.median(push) / 1e3:.2f} us, enter->exit "
                 f"{statistics.median(tot) / 1e3:.2f} us (max {max(tot) / 1e3:.2f})")
print(line, flush=True)
dist.barrier(); dist.destroy_process_group()
