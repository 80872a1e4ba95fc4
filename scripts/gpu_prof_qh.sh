#!/bin/bash
# ncu --set full of the [q;h] kernel (short bench command run plainly first)
mkdir -p gpurun_out
SHORT="python bench.py --steps 2 --warmup 3 --batch 8192 --e2e-batch 512 --no-cpu-baseline --no-lti"
$SHORT > gpurun_out/short_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'qh_kernel' -s 3 -c 1 \
    -f -o gpurun_out/prof_qh $SHORT > gpurun_out/full_ncu_qh.log 2>&1
echo "ncu exit $?"
