cd guided-extended-hybrid-astar_b200/csrc
for K in 5 3 4 6 8; do
  touch ga_potential.cu; make EXTRA=-DDT_K=$K > /dev/null 2>&1 || { echo "build failed K=$K"; continue; }
  cd ../..
  r=$(timeout 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_ref_golden.py -m gpu -x -q -k "potential or distance or dt" 2>&1 | tail -1)
  python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-plan --no-replan --no-collision --batch 0 > gpurun_out/b_k.json 2> gpurun_out/b_k.err
  python -c "
import json; d=json.load(open('gpurun_out/b_k.json')); print('K=$K', d['kernels_ms_per_step']['distance_transforms'], d['value'], '$r')"
  cd guided-extended-hybrid-astar_b200/csrc
done
