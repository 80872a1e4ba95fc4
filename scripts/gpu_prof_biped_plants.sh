#!/bin/bash
# ncu --set full of the two kernels of the embedded path (biped with per-instance plants),
# small batch; summaries + hottest source lines written on the box.
#   SPECS="couple:0 assemble:0" (kernel-name regex : launches to skip)
mkdir -p gpurun_out
CMD="python scripts/bench_biped_plants.py --batch ${B:-2368} --steps 1"
$CMD > gpurun_out/biped_plants_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/biped_plants_plain.log; exit 1; }
for spec in ${SPECS:-couple:0 assemble:0}; do
  k=${spec%%:*}; skip=${spec##*:}
  rep=gpurun_out/prof_bp_${k}
  ncu --set full --clock-control none --import-source on -k regex:lti_$k -s $skip -c 1 -f -o $rep $CMD > gpurun_out/full_ncu_bp_$k.log 2>&1
  echo "ncu $k exit $?"
  python scripts/ncu_summary.py $rep.ncu-rep > gpurun_out/ncu_bp_$k.txt 2>&1
  python scripts/ncu_lines.py $rep.ncu-rep 40 >> gpurun_out/ncu_bp_$k.txt 2>&1
  rm -f $rep.ncu-rep
done
