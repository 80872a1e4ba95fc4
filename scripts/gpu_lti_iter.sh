#!/bin/bash
# One iteration on the per-instance-plant kernel: parity, timings, optional ncu captures.
#   scripts/gpu_lti_iter.sh [prof] [all-tests]
mkdir -p gpurun_out
if [[ " $* " == *" all-tests "* ]]; then
  python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/pytest_gpu.log
else
  python -m pytest tests/test_gpu_lti.py tests/test_gpu_reference_live.py tests/test_gpu_guards.py -q -k "lti or plants or guard or lip or named or permuted" > gpurun_out/pytest_lti.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/pytest_lti.log
fi
python scripts/bench_lti.py --config ${CFG:-all} > gpurun_out/bench_lti.jsonl 2> gpurun_out/bench_lti.err; echo "bench_lti exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/bench_lti.jsonl'):
    d=json.loads(l)
    print(d['config'], d['hessian'], 'qps %.0f'%d['qps'], 'ms', {k: round(v,3) for k,v in d['ms'].items()}, 'GB/s all %.0f P %.0f Gh %.0f'%(d['hbm_gbs_all'], d['hbm_gbs_P'], d['hbm_gbs_Gh']), 'TF %.1f'%d['hessian_tflops_executed'], d['launch'])
PY
if [[ " $* " == *" prof "* ]]; then CFGS=c4 LAUNCHES="${LAUNCHES:-11 13 17 19}" bash scripts/gpu_prof_lti.sh > gpurun_out/prof_lti.log 2>&1; tail -5 gpurun_out/prof_lti.log; fi
