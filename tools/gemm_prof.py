"""One GEMM shape, a few launches -- the target of `ncu --set full` captures."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.ops import gemm

M, N, K, a_mn, b_mn, bn = (int(x) for x in sys.argv[1:7])
A = (torch.randn((K, M) if a_mn else (M, K), device="cuda") * 0.1).bfloat16()
B = (torch.randn((K, N) if b_mn else (N, K), device="cuda") * 0.1).bfloat16()
C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(5):
    gemm.gemm_bf16(A, B, C, a_mn=a_mn, b_mn=b_mn, bn=bn)
torch.cuda.synchronize()
print("ok", float(C.float().abs().mean()))
