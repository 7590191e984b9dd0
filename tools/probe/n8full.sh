nproc
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --steps 200 --warmup 5 > gpurun_out/full_n8.json 2> gpurun_out/full_n8.err
echo rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/full_n8_ref.json 2> gpurun_out/full_n8_ref.err
echo rc=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/full_n8.json') if l.startswith('{')][-1])
print(d['value'], d['e2e']['value'], d['e2e']['planes_value'], d['e2e']['lean_value'], d['e2e']['pipeline'], d['batch']['grids_per_s_all_ranks'], d['batch_large']['grids_per_s'], d['plan'].get('queries_per_s_all_ranks'), d['clocks'])
r=json.loads([l for l in open('gpurun_out/full_n8_ref.json') if l.startswith('{')][-1]); print(r['value'], r['n_gpus'], r['cpu_baseline']['cores'])"
tail -3 gpurun_out/full_n8.err
