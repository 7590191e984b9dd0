nproc
for t in 16 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 100 --warmup 5 --no-cpu-baseline --no-plan --no-replan --no-collision --batch 0 --e2e-threads $t > gpurun_out/n4_t$t.json 2> gpurun_out/n4_t$t.err
python -c "
import json
d=json.loads([l for l in open('gpurun_out/n4_t$t.json') if l.startswith('{')][-1]); e=d['e2e']; print($t, d['value'], e['value'], e['ms_per_step'], e.get('lean_value'), e['pipeline'])"
done
