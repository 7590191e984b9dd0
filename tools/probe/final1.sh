python -m pytest tests/ -x -q -m gpu 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err
python bench.py > gpurun_out/final_n1.json 2> gpurun_out/final_n1.err
python -c "
import json; d=json.load(open('gpurun_out/final_n1.json')); r=json.load(open('gpurun_out/final_ref.json'))
print({k:d[k] for k in ['value','ms_per_step','gpu_launches']}, d['kernels_ms_per_step'])
e=d['e2e']; print('e2e', e['value'], e['planes_value'], e['lean_value'], e['serial_value'], e['host_sweep_ms_per_step'])
print('roof', d['roofline']['frac'], d['safety_roofline']['frac'], 'batch', d['batch']['grids_per_s'], d['batch']['sweep_kernel_roofline_frac'], d['batch_large']['grids_per_s'], d['batch_large']['sweep_kernel_roofline_frac'])
print('plan', d['plan']['b200']['queries_per_s'], d['plan']['cpu_reference']['queries_per_s'], 'replan', d['replan_cycle']['ms_per_cycle'], 'coll', d['collision']['pose_checks_per_s'])
print('cpu1', d['cpu_baseline']['value'], 'ref arm', r['value'], r['cpu_baseline']['cores'], d['clocks'])"
