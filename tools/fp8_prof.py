"""One MX-fp8 GEMM shape, a few launches -- the target of `ncu --set full` captures."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_distributed_machine_learning_b200.ops.fp8 import gemm_mxfp8, quantize_mxfp8

M, N, K = (int(x) for x in sys.argv[1:4])
tile = sys.argv[4] if len(sys.argv) > 4 else "2cta192"
A = (torch.randn(M, K, device="cuda") * 0.1).bfloat16()
B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
qa, qb = quantize_mxfp8(A), quantize_mxfp8(B)
for _ in range(5):
    gemm_mxfp8(qa, qb, C, tile=tile)
torch.cuda.synchronize()
print("ok", float(C.float().abs().mean()))
