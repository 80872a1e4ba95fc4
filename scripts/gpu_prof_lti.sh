#!/bin/bash
# ncu --set full of the per-instance kernel (C5, C4; small batches).  bench_lti.py launches,
# per mode, [all, all(x reps), P, P(x reps), q.., Gh.., none..]; with --reps 1 the kernel
# launches are: DMMA 0-9 (all,all,P,P,q,q,Gh,Gh,0,0), recursion 10-19.
# Captured: launch 3 (DMMA P-only), 13 (recursion P-only), 17 (recursion G,h-only), 11 (recursion all).
mkdir -p gpurun_out
for cfg in ${CFGS:-c5 c4}; do
  if [ $cfg = c5 ]; then B=296; else B=2368; fi
  CMD="python scripts/bench_lti.py --config $cfg --batch $B --reps 1"
  $CMD > gpurun_out/lti_plain_$cfg.log 2>&1 || { echo "plain run failed"; exit 1; }
  for skip in ${LAUNCHES:-3 13 17 11}; do
    ncu --set full --clock-control none --import-source on -k regex:lti_assemble -s $skip -c 1 \
        -f -o gpurun_out/prof_lti_${cfg}_l$skip $CMD > gpurun_out/full_ncu_lti_${cfg}_l$skip.log 2>&1
    echo "ncu $cfg launch $skip exit $?"
  done
done
ls -la gpurun_out
