timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "staging_modes or threshold_edges or safety or dilation or empty" 2>&1 | tail -4
timeout 300 bash tools/probe/quick.sh
