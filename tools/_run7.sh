mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/r2_gpu_tests_n1.log 2>&1; echo rc=$?; tail -12 gpurun_out/r2_gpu_tests_n1.log
python - <<'PY' 2>&1 | tail -5
import torch, sys
sys.path.insert(0,'.')
from simple_distributed_machine_learning_b200.ops.loss import softmax_ce
x=(torch.randn(2048,8192,device='cuda')*2).bfloat16(); t=torch.randint(0,8192,(2048,),device='cuda')
d=torch.empty_like(x); ls=torch.zeros(1,device='cuda')
flush=torch.empty(256<<20,dtype=torch.uint8,device='cuda')
ts=[]
for i in range(12):
    flush.zero_()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record(); softmax_ce(x,t,grad_scale=1.0,loss_sum=ls,dlogits=d); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ms=sorted(ts)[len(ts)//2]
print(f"softmax_ce 2048x8192 bf16 (L2 flushed): {ms*1e3:.1f} us, {2*x.numel()*2/ms/1e6:.0f} GB/s of 6536 measured copy bandwidth")
PY
