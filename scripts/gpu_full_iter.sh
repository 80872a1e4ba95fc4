#!/bin/bash
# Full GPU parity suite + the per-instance-plant timings (+ optional extras / K1 bench).
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_gpu.log
python scripts/bench_lti.py --config all > gpurun_out/bench_lti.jsonl 2> gpurun_out/bench_lti.err; echo "bench_lti exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/bench_lti.jsonl'):
    d=json.loads(l)
    print(d['config'], d['hessian'], 'qps %.0f'%d['qps'], 'ms', {k: round(v,3) for k,v in d['ms'].items()}, 'GB/s all %.0f P %.0f Gh %.0f'%(d['hbm_gbs_all'], d['hbm_gbs_P'], d['hbm_gbs_Gh']), 'TF %.1f'%d['hessian_tflops_executed'], d['launch'])
PY
python scripts/bench_extend.py > gpurun_out/bench_extend.jsonl 2> gpurun_out/bench_extend.err; echo "bench_extend exit $?"; cat gpurun_out/bench_extend.jsonl | cut -c1-300
python scripts/bench_extras.py > gpurun_out/bench_extras.json 2> gpurun_out/bench_extras.err; echo "extras exit $?"; cat gpurun_out/bench_extras.json | head -60
