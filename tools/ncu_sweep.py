"""One launch of every kernel on a benchmarked path inside a cudaProfilerStart/Stop range, for

    ncu --set full --clock-control none --import-source on --profile-from-start off \
        --kernel-name regex:sdml -o gpurun_out/sweep_r2 python tools/ncu_sweep.py

(one GPU; tools/make_evidence.py turns the report into profiles/ncu_sweep_r2.md).  The collective
kernels run here with a single rank -- ncu must not wrap a multi-rank job -- so their NVLink
traffic is not in this capture; their cross-GPU timing comes from the in-kernel globaltimer probes
(profiles/dp_ll_kernel_n*_r2.log) and the multi-GPU bench lines."""
import ctypes
import io
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.models.mlp import MLPStage  # noqa: E402
from simple_distributed_machine_learning_b200.ops import _lib  # noqa: E402
from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine  # noqa: E402
from simple_distributed_machine_learning_b200.parallel.symm import FusedDPSGD  # noqa: E402
from simple_distributed_machine_learning_b200.parallel.topology import Topology  # noqa: E402
from simple_distributed_machine_learning_b200.train import Trainer  # noqa: E402
from simple_distributed_machine_learning_b200.utils.config import TrainConfig  # noqa: E402

dev = torch.device("cuda", 0)
lib = _lib.cuda_lib()
H, B = 8192, 2048

# ---- set-up + warm-up (not profiled) ----
engines = []
for dtype in ("bf16", "fp8"):
    st = MLPStage(0, 1, 2, H, H, dev, lr=0.01, momentum=0.5, dtype=dtype)
    engines.append(PipelineEngine(st, Topology(1, 1), 0, None, microbatches=1, schedule="gpipe"))
x = torch.randn(B, H, device=dev).bfloat16()
y = torch.randint(0, H, (B,), device=dev)
for e in engines:
    for _ in range(2):
        e.train_step(x, y, B)
cfg = TrainConfig(rank=0, world_size=1, device="cuda", epochs=1, train_samples=240, test_samples=60, use_cuda_graph=False)
tr = Trainer(cfg, out=io.StringIO())
lx = torch.randn(60, 1, 28, 28, device=dev)
ly = torch.randint(0, 10, (60,), device=dev)
for _ in range(2):
    tr.engine.train_step(lx, ly, 60)
shapes = [(10, 1, 5, 5), (10,), (20, 10, 5, 5), (20,), (50, 320), (50,), (10, 50), (10,)]
ps = [torch.randn(s, device=dev) for s in shapes]
gs = [torch.randn(s, device=dev) for s in shapes]
ms = [torch.zeros(s, device=dev) for s in shapes]
dp = FusedDPSGD(ps, lambda: gs, ms, 0.1, 0.5, 0.0, [0], 0, None, dev)
dp.step()
src = torch.randn(8 * 320, device=dev)
dst = torch.empty_like(src)
words = torch.zeros(16, dtype=torch.int32, device=dev)     # [0] flag, [1] counter, [4..7] error record
s0 = torch.cuda.current_stream().cuda_stream
P = ctypes.c_void_p


def boundary_once():
    lib.sdml_bnd_push(P(dst.data_ptr()), P(src.data_ptr()), src.numel() * 4, P(words.data_ptr()),
                      P(words.data_ptr() + 4), 1, None, 0, P(words.data_ptr() + 16), None, 3, s0)
    lib.sdml_bnd_wait(P(words.data_ptr()), P(words.data_ptr() + 4), 1, 10 ** 9, P(words.data_ptr() + 16), None, 1, s0)
    lib.sdml_bnd_signal(P(words.data_ptr()), P(words.data_ptr() + 4), 1, 1, s0)


boundary_once()
torch.cuda.synchronize()

# ---- the profiled range: every kernel once (a training step of each family) ----
torch.cuda.profiler.start()
for e in engines:
    e.train_step(x, y, B)
tr.engine.train_step(lx, ly, 60)
dp.step()
boundary_once()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ncu sweep range done")
