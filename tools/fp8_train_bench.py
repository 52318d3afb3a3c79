"""Step time of the wide MLP on one GPU: bf16 vs fp8 forward vs fp8 forward+backward."""
import os
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.models.mlp import MLPStage
from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
from simple_distributed_machine_learning_b200.parallel.topology import Topology

H = int(os.environ.get("H", 8192)); L = int(os.environ.get("L", 8)); B = int(os.environ.get("B", 16384))
MB = int(os.environ.get("MB", 8))
x = torch.randn(B, H, device="cuda").bfloat16()
t = torch.randint(0, H, (B,), device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name, dtype, bwd in (("bf16", "bf16", "0"), ("fp8-fwd", "fp8", "0"), ("fp8-all", "fp8", "1")):
    os.environ["SDML_FP8_BWD"] = bwd
    st = MLPStage(0, 1, L, H, H, torch.device("cuda"), lr=0.01, momentum=0.5, dtype=dtype)
    eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=MB, schedule="gpipe")
    losses = [eng.train_step(x, t, B) for _ in range(3)]
    torch.cuda.synchronize()
    ts = []
    for _ in range(6):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); loss = eng.train_step(x, t, B); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    fl = st.flops_per_sample() * B
    print(f"{name}: {ms:.2f} ms/step, {B / ms * 1e3:.0f} samples/s, {fl / ms / 1e9:.0f} TFLOP/s model, "
          f"loss {losses[0].item():.4f} -> {loss.item():.4f}", flush=True)
    del st, eng
    torch.cuda.empty_cache()
