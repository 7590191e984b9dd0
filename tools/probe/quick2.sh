python -m pytest tests -m gpu -x -q 2>&1 | tail -2; python tools/knob_sweep.py | tail -1
