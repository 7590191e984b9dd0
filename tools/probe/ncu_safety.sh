CMD="python bench.py --steps 6 --warmup 3 --batch 0 --no-cpu-baseline --no-collision --no-replan --no-plan --e2e-threads 4"
ncu --set full --clock-control none --import-source on -k regex:safety_kernel -s 4 -c 1 -f -o gpurun_out/full_safety_v4 $CMD > gpurun_out/ncu_safety_v4.log 2>&1
tail -3 gpurun_out/ncu_safety_v4.log
