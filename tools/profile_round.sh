#!/bin/bash
# Round profile capture (run on the GPU box through gpurun): unprofiled bench first, then the ncu
# launch list and one --set full capture per dominant kernel.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 6 --warmup 3 --batch 0 --no-cpu-baseline --no-collision --no-replan"
$CMD > gpurun_out/prof_bench.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/prof_bench.log; exit 1; }
tail -1 gpurun_out/prof_bench.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
for k in heur_sweep_kernel safety_kernel heur_finish_a_kernel heur_finish_b_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/full_$k $CMD > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out
