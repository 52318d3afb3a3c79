"""Small-shape run of every kernel family, meant to be executed under
`compute-sanitizer --tool memcheck` (one tool per gpurun call)."""
import io, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.ops import gemm
from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8
from simple_distributed_machine_learning_b200.ops.loss import softmax_ce
from simple_distributed_machine_learning_b200.models.mlp import MLPStage
from simple_distributed_machine_learning_b200.parallel.pipeline import PipelineEngine
from simple_distributed_machine_learning_b200.parallel.topology import Topology
from simple_distributed_machine_learning_b200.train import Trainer
from simple_distributed_machine_learning_b200.utils.config import TrainConfig

dev = torch.device("cuda")
for (M, N, K, a_mn, b_mn, bn) in ((200, 136, 192, 0, 0, 0), (256, 256, 128, 0, 1, 1256), (520, 264, 320, 1, 1, 1256),
                                  (512, 512, 256, 0, 0, 1128), (136, 72, 64, 1, 0, 64)):
    A = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
    B = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
    for out_fp32 in (False, True):
        C = torch.zeros(M, N, device=dev, dtype=torch.float32 if out_fp32 else torch.bfloat16)
        gemm.gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bias=torch.randn(N, device=dev), relu=True,
                       accumulate=out_fp32, bn=bn)
qa, qb = quantize_mxfp8(torch.randn(200, 256, device=dev).bfloat16()), quantize_mxfp8(torch.randn(136, 256, device=dev).bfloat16())
gemm_mxfp8(qa, qb, torch.empty(200, 136, device=dev, dtype=torch.bfloat16), relu=True)
softmax_ce(torch.randn(33, 1000, device=dev).bfloat16(), torch.randint(0, 1000, (33,), device=dev), want_correct=True)
st = MLPStage(0, 1, 2, 256, 128, dev, lr=0.05, momentum=0.5)
eng = PipelineEngine(st, Topology(1, 1), 0, None, microbatches=2, schedule="zb-h1")
x, y = torch.randn(512, 256, device=dev), torch.randint(0, 128, (512,), device=dev)
for _ in range(2):
    eng.train_step(x, y, 512)
t = Trainer(TrainConfig(rank=0, world_size=1, device="cuda", epochs=1, train_samples=120, test_samples=60,
                        use_cuda_graph=False), out=io.StringIO())
t.fit()
torch.cuda.synchronize()
print("sanitize run complete")
