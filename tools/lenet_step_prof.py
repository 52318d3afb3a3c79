"""A few reference-model (LeNet, batch 60) training steps on one GPU, eager (no CUDA graph),
as the target of an ncu launch list."""
import io, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.train import Trainer
from simple_distributed_machine_learning_b200.utils.config import TrainConfig

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
cfg = TrainConfig(rank=0, world_size=1, device="cuda", epochs=1, train_samples=600, test_samples=60,
                  use_cuda_graph=False)
t = Trainer(cfg, out=io.StringIO())
it = iter(t.train_loader)
for _ in range(steps):
    d, y = next(it)
    loss = t.train_batch(d, y)
torch.cuda.synchronize()
print("ok", float(loss))
