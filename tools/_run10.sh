mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for db in 1 0; do
SDML_E2E_DOUBLE_BUFFER=$db timeout 300 $TR --master-port 2968$db bench.py --gpus 2 --steps 200 --warmup 20 --no-extra --no-baseline > gpurun_out/r2_n2_db$db.json 2> gpurun_out/r2_n2_db$db.err; echo rc=$?
python -c "
import json; r=json.load(open('gpurun_out/r2_n2_db$db.json')); print('double_buffer=$db headline', round(r['value']), 'e2e', round(r['e2e']['value']), round(r['e2e']['ms_per_step']*1e3,1),'us')"
done
