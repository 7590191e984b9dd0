timeout 600 python -m pytest tests/test_gpu_plan_parity.py tests/test_gpu_round2.py -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-plan --no-replan --no-collision --batch 0 > gpurun_out/b_rec.json 2> gpurun_out/b_rec.err; tail -3 gpurun_out/b_rec.err
python -c "
import json; d=json.load(open('gpurun_out/b_rec.json')); e=d['e2e']; print(d['value'], e['value'], e['ms_per_step'], e['planes_value'], e['lean_value'], e['serial_value'], e['d2h_bytes_per_step'])"
