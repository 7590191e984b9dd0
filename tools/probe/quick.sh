python -m pytest tests -m gpu -x -q 2>&1 | tail -3; python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-plan --no-replan --no-collision --batch 128 > gpurun_out/b_s1.json 2> gpurun_out/b_s1.err; tail -3 gpurun_out/b_s1.err; python -c "
import json; d=json.load(open('gpurun_out/b_s1.json')); print(d['value'], d['kernels_ms_per_step'], d['safety_roofline']['frac'], d.get('batch'), d.get('batch_large'))"
python tools/batch_stats.py 128 2048 | tail -1
