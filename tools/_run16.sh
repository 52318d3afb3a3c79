mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29741 bench.py --gpus 2 --steps 20 --warmup 5 --extras lenet_pp,mlp,mlp_fp8 --no-baseline --no-bubble > gpurun_out/r2_bench_n2_e2e.json 2> gpurun_out/r2_bench_n2_e2e.err; echo rc=$?
python -c "
import json; r=json.load(open('gpurun_out/r2_bench_n2_e2e.json'))
print('headline', round(r['value']), 'e2e', round(r['e2e']['value']))
for k in ('lenet_pp','mlp','mlp_fp8'):
    v=r[k]; print(k, v.get('error') or (round(v['value']), 'ms/step', round(v['ms_per_step'],3), 'e2e', round(v['e2e']['value']), 'e2e ms', round(v['e2e']['ms_per_step'],3)))"
tail -2 gpurun_out/r2_bench_n2_e2e.err
