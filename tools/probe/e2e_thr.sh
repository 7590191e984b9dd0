nproc
for t in 8 12 16 24; do
python bench.py --steps 64 --warmup 5 --no-cpu-baseline --no-plan --no-replan --no-collision --batch 0 --e2e-threads $t > gpurun_out/b_thr$t.json 2> gpurun_out/b_thr$t.err; python -c "
import json; d=json.load(open('gpurun_out/b_thr$t.json')); e=d['e2e']; print($t, d['value'], e['value'], e['ms_per_step'], e['host_sweep_ms_per_step'], e.get('lean_value'))"
done
