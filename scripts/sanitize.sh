#!/bin/bash
# compute-sanitizer (race detection / sanitizers) over the smoke batch
# (mode A: stream_terms, qh_kernel, K1, the plan-build GEMMs) and a small per-instance-plant
# batch in both Hessian modes (mode B: lti_assemble_kernel with its bulk copies).
# Logs: gpurun_out/sanitizer_<tool>_<what>.log
mkdir -p gpurun_out
SMOKE="python __graft_entry__.py smoke"
LTI="python scripts/bench_lti.py --config small --batch 48 --reps 1"
for tool in memcheck racecheck; do
  for what in smoke lti; do
    if [ $what = smoke ]; then CMD=$SMOKE; else CMD=$LTI; fi
    timeout 400 compute-sanitizer --tool $tool --print-limit 20 $CMD > gpurun_out/sanitizer_${tool}_${what}.log 2>&1
    echo "$tool $what exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|smoke ok" gpurun_out/sanitizer_${tool}_${what}.log | tail -3
  done
done
