timeout 800 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_round2.py -m gpu -x -q -k "noalloc or records or near_nodes or download_window" 2>&1 | tail -8
echo "rc=$?"
timeout 800 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "staging_modes or threshold_edges or (safety_and_pool and not 4096 and not 2048 and not 2050) or potential" 2>&1 | tail -8
echo "rc=$?"
