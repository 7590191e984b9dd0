#!/bin/bash
# Round profile capture (run on the GPU box through gpurun): unprofiled bench first, then the ncu
# launch list and one --set full capture per dominant kernel.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 6 --warmup 3 --batch 0 --no-cpu-baseline --no-collision --no-replan --no-plan --e2e-threads 4"
$CMD > gpurun_out/prof_bench.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/prof_bench.log; exit 1; }
tail -1 gpurun_out/prof_bench.log | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
for k in heur_sweep_kernel safety_kernel dt_kernel heur_finish_a_kernel heur_finish_b_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/full_$k $CMD > gpurun_out/ncu_$k.log 2>&1
done
# the sweep kernel in throughput mode: one launch over 64 grids of 512^2 cells
BCMD="python tools/batch_stats.py 64 2048"
$BCMD > gpurun_out/prof_batch.log 2>&1 && tail -1 gpurun_out/prof_batch.log
ncu --set full --clock-control none --import-source on -k regex:heur_sweep_kernel -s 2 -c 1 -f -o gpurun_out/full_batch_heur_sweep_kernel $BCMD > gpurun_out/ncu_batch.log 2>&1
# collision path: one expansion fan, the asynchronous look-ahead launch, a large pose batch and an expansion batch
LCMD="python tools/latency.py 2048"
$LCMD > gpurun_out/prof_latency.log 2>&1 && tail -14 gpurun_out/prof_latency.log
for k in check_fan_kernel check_poses_kernel expand_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/full_$k $LCMD > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out | head -40
