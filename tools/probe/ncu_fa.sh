ncu --set full --clock-control none --import-source on -k regex:heur_finish_a_kernel -s 2 -c 1 -f -o gpurun_out/full_batch_finish_a python tools/batch_stats.py 128 2048 > gpurun_out/ncu_bfa.log 2>&1
tail -2 gpurun_out/ncu_bfa.log
