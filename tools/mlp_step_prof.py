"""A few wide-MLP training steps on one GPU (target of the ncu launch list)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.models.mlp import MLPStage
from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
from simple_distributed_machine_learning_b200.parallel.topology import Topology

H = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
MB = int(sys.argv[3]) if len(sys.argv) > 3 else 2048
M = int(sys.argv[4]) if len(sys.argv) > 4 else 8
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
dtype = sys.argv[6] if len(sys.argv) > 6 else "bf16"
d = torch.device("cuda")
st = MLPStage(0, 1, L, H, H, d, lr=0.01, momentum=0.5, dtype=dtype)
eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=M, schedule="gpipe")
x = torch.randn(MB * M, H, device=d).bfloat16()
t = torch.randint(0, H, (MB * M,), device=d)
for _ in range(steps):
    loss = eng.train_step(x, t, MB * M)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
loss = eng.train_step(x, t, MB * M)
e1.record()
torch.cuda.synchronize()
print("step ms", e0.elapsed_time(e1), "loss", float(loss))
