mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_fp8_gpu.py -m gpu -q --timeout 300 -k "tma_store" > gpurun_out/r2_fp8_tma_test.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2_fp8_tma_test.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29731 bench.py --gpus 2 --steps 20 --warmup 5 --extras lenet_pp,mlp,mlp_fp8 --no-baseline --no-bubble > gpurun_out/r2_bench_n2_fp8tma.json 2> gpurun_out/r2_bench_n2_fp8tma.err; echo rc=$?
python -c "
import json; r=json.load(open('gpurun_out/r2_bench_n2_fp8tma.json'))
for k in ('lenet_pp','mlp','mlp_fp8'):
    v=r[k]; print(k, v.get('error') or (round(v['value']), 'ms/step', round(v['ms_per_step'],3), 'e2e', round(v['e2e']['value'])))"
tail -2 gpurun_out/r2_bench_n2_fp8tma.err
