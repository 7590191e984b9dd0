timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q -k "heuristic or heur" 2>&1 | tail -2
python tools/batch_stats.py 128 2048 | tail -1
ncu --set full --clock-control none --import-source on -k regex:heur_finish_a_kernel -s 4 -c 1 -f -o gpurun_out/full_heur_finish_a_kernel python bench.py --steps 6 --warmup 3 --batch 0 --no-cpu-baseline --no-collision --no-replan --no-plan --e2e-threads 4 > gpurun_out/ncu_heur_finish_a_kernel.log 2>&1
