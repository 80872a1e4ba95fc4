#!/bin/bash
# ncu --set full of the scheduled-cost stage (lti_sched_kernel): launch SKIP of the bench below
# (the launches alternate biped / generic workloads; 0.. = biped with terminal cost).
mkdir -p gpurun_out
CMD="python scripts/bench_sched_costs.py --batch ${B:-4736} --steps 1"
$CMD > gpurun_out/sched_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sched_plain.log; exit 1; }
for skip in ${SKIPS:-0 5}; do
  rep=gpurun_out/prof_sched_$skip
  ncu --set full --clock-control none --import-source on -k regex:lti_sched -s $skip -c 1 -f -o $rep $CMD > gpurun_out/full_ncu_sched_$skip.log 2>&1
  echo "ncu skip $skip exit $?"
  python scripts/ncu_summary.py $rep.ncu-rep > gpurun_out/ncu_sched_$skip.txt 2>&1
  python scripts/ncu_lines.py $rep.ncu-rep 30 >> gpurun_out/ncu_sched_$skip.txt 2>&1
  rm -f $rep.ncu-rep
done
