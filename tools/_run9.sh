mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 300 -k "symmetric_boundary" > gpurun_out/r2_two_stream_tests_n2.log 2>&1; echo rc=$?; tail -5 gpurun_out/r2_two_stream_tests_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29671 bench.py --gpus 2 --steps 20 --warmup 5 --extras lenet_pp --no-baseline --no-bubble > gpurun_out/r2_bench_n2_lenetpp2s.json 2> gpurun_out/r2_bench_n2_lenetpp2s.err; echo rc=$?
python -c "
import json; r=json.load(open('gpurun_out/r2_bench_n2_lenetpp2s.json')); v=r['lenet_pp']; print('lenet_pp', round(v['value']), 'us/step', round(v['ms_per_step']*1e3,1), 'e2e', round(v['e2e']['value'])); print('headline', round(r['value']), 'e2e', round(r['e2e']['value']))"
tail -3 gpurun_out/r2_bench_n2_lenetpp2s.err
