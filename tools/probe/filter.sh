timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "heuristic" 2>&1 | tail -3
python tools/batch_stats.py 128 2048 | tail -1
python tools/batch_stats.py 128 2048 2 | tail -1
python tools/heur_stats.py 2>&1 | tail -3
