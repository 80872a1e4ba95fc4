#!/bin/bash
# Rebuilds mpci_lti.cu on the GPU box with different -D settings and times each variant
# with scripts/bench_lti.py:   VARIANTS="-DMPCI_ISSUE_LANES=1 -DMPCI_ISSUE_LANES=4" bash scripts/gpu_variants.sh
mkdir -p gpurun_out
cd mpc_interface_b200/csrc
for v in $VARIANTS; do
  nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -I../../include -I. $(echo $v | tr ',' ' ') -c mpci_lti.cu -o mpci_lti.o || exit 1
  nvcc -shared -gencode arch=compute_100a,code=sm_100a -o libmpci.so mpci_api.o mpci_extend.o mpci_linalg.o mpci_assemble.o mpci_lti.o -lcudart || exit 1
  for cfg in ${CFGS:-c4 c5}; do
    (cd ../..; python scripts/bench_lti.py --config $cfg | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if d['hessian'] == 'recursion':
        print('$v', d['config'], 'qps %.0f' % d['qps'], {k: round(x, 3) for k, x in d['ms'].items()})
")
  done
done 2>&1 | tee ../../gpurun_out/variants.log
