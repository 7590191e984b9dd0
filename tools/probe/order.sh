timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "batch_matches" 2>&1 | tail -3
for cfg in "-1" "0 50 1" "0 50 0" "16 50 1" "16 50 0" "16 25 0" "16 100 0" "32 25 0" "32 12 0" "16 200 0" "8 100 0"; do
  timeout 120 python tools/batch_stats.py 128 2048 $cfg 2>&1 | tail -1
done
