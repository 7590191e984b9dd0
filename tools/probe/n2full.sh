python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 5 > gpurun_out/full_n2.json 2> gpurun_out/full_n2.err
echo rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/full_n2_ref.json 2> gpurun_out/full_n2_ref.err
echo rc=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/full_n2.json') if l.startswith('{')][-1])
print(d['value'], d['e2e']['value'], d['e2e']['pipeline'], d['batch']['grids_per_s_all_ranks'], d['batch_large']['grids_per_s'], d['plan'].get('queries_per_s_all_ranks'), d['clocks'])
r=json.loads([l for l in open('gpurun_out/full_n2_ref.json') if l.startswith('{')][-1]); print(r['value'], r['n_gpus'], r['cpu_baseline']['cores'])"
