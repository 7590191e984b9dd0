timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q -k "heuristic or heur" 2>&1 | tail -2
timeout 200 python tools/knob_sweep.py 8,20,-1,-1,-1,-1,5 8,20,-1,-1,-1,-1,1 8,20,-1,-1,-1,-1,9 8,20,-1,-1,-1,-1,1 8,20,-1,-1,-1,-1,9 2>&1 | tail -6
