"""Warm, per-kernel device times of the fused LeNet step (CUDA events, 200 iterations each)."""
import io, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.train import Trainer
from simple_distributed_machine_learning_b200.utils.config import TrainConfig

cfg = TrainConfig(rank=0, world_size=1, device="cuda", epochs=1, train_samples=600, test_samples=60,
                  use_cuda_graph=False)
t = Trainer(cfg, out=io.StringIO())
st = t.stage
d, y = next(iter(t.train_loader))
x, tt = d.cuda(), y.cuda()
st.begin_step(60)


def timeit(fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / n


s = st._conv_fwd(x, 0, 0.5)
print("s0_fwd   %.2f us" % timeit(lambda: st._conv_fwd(x, 0, 0.5)))
print("s1(+wg)  %.2f us" % timeit(lambda: st._fc(s["z3"], tt, 0, True, 0.5)))
dz3 = st._fc(s["z3"], tt, 0, True, 0.5)
print("s0_bwd   %.2f us" % timeit(lambda: st._conv_bwd(s, dz3)))
print("sgd      %.2f us" % timeit(lambda: st.optimizer.step(zero_grad=True)))
print("step eager %.2f us" % timeit(lambda: t.engine.train_step(x, tt, 60), 100))
t2 = Trainer(TrainConfig(rank=0, world_size=1, device="cuda", epochs=1, train_samples=600, test_samples=60),
             out=io.StringIO())
t2.train_step_device(x, tt, 60)
print("step graph %.2f us" % timeit(lambda: t2.train_step_device(x, tt, 60), 300))
