mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_tp_gpu.py -m gpu -q --timeout 300 > gpurun_out/r2_tp_tests_n2.log 2>&1; echo rc=$?; tail -4 gpurun_out/r2_tp_tests_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 500 $TR --master-port 29751 bench.py --gpus 2 --steps 20 --warmup 5 --extras mlp_tp_fp8 --mlp-steps 10 --no-baseline --no-bubble > gpurun_out/r2_bench_n2_tpfp8.json 2> gpurun_out/r2_bench_n2_tpfp8.err; echo rc=$?
python -c "
import json; r=json.load(open('gpurun_out/r2_bench_n2_tpfp8.json'))
for k in ('mlp_tp_fp8',):
    v=r[k]; print(k, v.get('error') or (round(v['value']), 'ms/step', round(v['ms_per_step'],3), 'e2e', round(v['e2e']['value']), v.get('roofline')))"
grep -v "unbatched" gpurun_out/r2_bench_n2_tpfp8.err | tail -3
