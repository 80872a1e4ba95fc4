#!/bin/bash
# One GPU-box pass: parity tests, smoke, measured peaks, bench.  Outputs in gpurun_out/.
#   scripts/gpu_check.sh [notests] [noprof]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python -c "import os; print('cpus', os.cpu_count(), len(os.sched_getaffinity(0)))" >> gpurun_out/gpu.txt
free -g >> gpurun_out/gpu.txt
if [[ " $* " != *" notests "* ]]; then
  python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
  tail -25 gpurun_out/pytest_gpu.log
fi
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
scripts/peaks.bin > gpurun_out/peaks.json 2> gpurun_out/peaks.err; cat gpurun_out/peaks.json; tail -2 gpurun_out/peaks.err
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -c 6000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
if [[ " $* " == *" noprof "* ]]; then exit 0; fi
SHORT="python bench.py --steps 2 --warmup 3 --batch 8192 --no-e2e --no-cpu-baseline --no-lti"
$SHORT > gpurun_out/short_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches.csv $SHORT > gpurun_out/short_ncu.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:'stream_terms|qh_kernel' -s 9 -c 3 \
    -f -o gpurun_out/prof_assemble $SHORT > gpurun_out/full_ncu.log 2>&1
echo "ncu full exit $?"
