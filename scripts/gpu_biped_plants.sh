#!/bin/bash
# One iteration on the embedded per-instance path (biped with one plant per instance):
# parity against the live reference, the plain per-instance tests, timings.
mkdir -p gpurun_out
[[ " $* " == *" notests "* ]] || python -m pytest tests/test_gpu_biped_plants.py -q -x > gpurun_out/pytest_biped_plants.log 2>&1; echo "pytest biped plants exit $?"; tail -25 gpurun_out/pytest_biped_plants.log
if [[ " $* " == *" lti "* ]]; then
  python -m pytest tests/test_gpu_lti.py tests/test_gpu_reference_live.py tests/test_gpu_guards.py -q -k "lti or plants or guard or lip or named or permuted" > gpurun_out/pytest_lti.log 2>&1; echo "pytest lti exit $?"; tail -5 gpurun_out/pytest_lti.log
  python scripts/bench_lti.py --config ${CFG:-all} > gpurun_out/bench_lti.jsonl 2> gpurun_out/bench_lti.err; echo "bench_lti exit $?"
  python - <<'PY'
import json
for l in open('gpurun_out/bench_lti.jsonl'):
    d=json.loads(l)
    print(d['config'], d['hessian'], 'qps %.0f'%d['qps'], 'ms', {k: round(v,3) for k,v in d['ms'].items()}, 'GB/s all %.0f'%d['hbm_gbs_all'], 'TF %.1f'%d['hessian_tflops_executed'])
PY
fi
if [ -f scripts/bench_biped_plants.py ]; then python scripts/bench_biped_plants.py > gpurun_out/bench_biped_plants.json 2> gpurun_out/bench_biped_plants.err; echo "bench biped plants exit $?"; cat gpurun_out/bench_biped_plants.json; tail -3 gpurun_out/bench_biped_plants.err; fi
