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
if [[ " $* " == *" prof "* ]]; then
  # launch list of the two kernels; bench_biped_plants launches, per mode, [all, P, q, Gh] x (3 + steps)
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:lti_ -c 400 --csv --log-file gpurun_out/launches_biped_plants.csv python scripts/bench_biped_plants.py --steps 2 > gpurun_out/ncu_biped_plants.log 2>&1
  python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/launches_biped_plants.csv')) if len(r) > 5]
head = rows[0]; ki, vi = head.index('Kernel Name'), head.index('Metric Value')
seq = collections.OrderedDict()
for r in rows[1:]:
    try: v = float(r[vi].replace(',', ''))
    except ValueError: continue
    seq.setdefault(r[ki][:60], []).append(v / 1e6)
for k, v in seq.items():
    per = 5
    print(k, len(v), "launches;", "ms per block of %d:" % per, [round(sum(v[i:i + per]) / len(v[i:i + per]), 3) for i in range(0, len(v), per)])
PY
fi
