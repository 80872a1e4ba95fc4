#!/bin/bash
# ncu --set full of the per-instance kernel (C5, C4; small batches): the P-only launch in
# DMMA mode (launch 2), the G,h-only launch (6) and the P-only launch in recursion mode (12).
mkdir -p gpurun_out
for cfg in c5 c4; do
  if [ $cfg = c5 ]; then B=296; else B=2368; fi
  CMD="python scripts/bench_lti.py --config $cfg --batch $B --reps 1"
  $CMD > gpurun_out/lti_plain_$cfg.log 2>&1 || { echo "plain run failed"; exit 1; }
  for skip in 2 6 12; do
    ncu --set full --clock-control none --import-source on -k regex:lti_assemble -s $skip -c 1 \
        -f -o gpurun_out/prof_lti_${cfg}_l$skip $CMD > gpurun_out/full_ncu_lti_${cfg}_l$skip.log 2>&1
    echo "ncu $cfg launch $skip exit $?"
  done
done
ls -la gpurun_out
