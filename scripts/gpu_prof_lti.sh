#!/bin/bash
# ncu --set full of the per-instance kernel (small batches).  bench_lti.py launches, per mode,
# [all, all(x reps), P, P(x reps), q.., Gh.., none..]; with --reps 1 the kernel launches are:
# DMMA 0-9 (all,all,P,P,q,q,Gh,Gh,0,0), recursion 10-19.
# Captured: launch 3 (DMMA P-only), 13 (recursion P-only), 17 (recursion G,h-only), 11 (recursion all),
# 19 (prediction only).  The reports are summarised on the box (metrics + hottest source lines);
# KEEP="c4_l11 ..." keeps those .ncu-rep files (gpurun brings back at most 64 MiB).
mkdir -p gpurun_out
for cfg in ${CFGS:-c5 c4}; do
  if [ $cfg = c5 ]; then B=296; else B=2368; fi
  CMD="python scripts/bench_lti.py --config $cfg --batch $B --reps 1"
  $CMD > gpurun_out/lti_plain_$cfg.log 2>&1 || { echo "plain run failed"; exit 1; }
  for skip in ${LAUNCHES:-3 13 17 11 19}; do
    rep=gpurun_out/prof_lti_${cfg}_l$skip
    ncu --set full --clock-control none --import-source on -k regex:lti_assemble -s $skip -c 1 \
        -f -o $rep $CMD > gpurun_out/full_ncu_lti_${cfg}_l$skip.log 2>&1
    echo "ncu $cfg launch $skip exit $?"
    python scripts/ncu_summary.py $rep.ncu-rep > gpurun_out/ncu_lti_${cfg}_l$skip.txt 2>&1
    python scripts/ncu_lines.py $rep.ncu-rep 25 >> gpurun_out/ncu_lti_${cfg}_l$skip.txt 2>&1
    case " ${KEEP:-} " in *" ${cfg}_l$skip "*) ;; *) rm -f $rep.ncu-rep ;; esac
  done
done
ls -la gpurun_out | head -40
