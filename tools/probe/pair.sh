timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q -k "heuristic or heur" 2>&1 | tail -3
timeout 200 python tools/knob_sweep.py 8,20,-1,-1,-1,-1,5 6,20,-1,-1,-1,-1,1 4,20,-1,-1,-1,-1,1 2>&1 | tail -4
