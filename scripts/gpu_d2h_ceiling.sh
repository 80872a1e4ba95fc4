#!/bin/bash
# Plain-copy device -> pinned-host ceiling of the box with N GPUs copying at once, N = 1, 2, 4, 8
# (one cudaMemcpyAsync stream of 1 GiB pieces per GPU, scripts/peaks.cu --d2h-only): what the
# host-buffer entry points are up against.  Then the c3 bench line on all GPUs of the box.
#   gpurun --gpus 8 -- bash scripts/gpu_d2h_ceiling.sh
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "[" > gpurun_out/d2h_ceiling.json
first=1
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  for i in $(seq 0 $((n - 1))); do
    CUDA_VISIBLE_DEVICES=$i scripts/peaks.bin --d2h-only 12 > gpurun_out/d2h_${n}_$i.json 2> gpurun_out/d2h_${n}_$i.err &
  done
  wait
  python - $n >> gpurun_out/d2h_ceiling.json <<'PY'
import json, sys
n = int(sys.argv[1])
per = [json.load(open('gpurun_out/d2h_%d_%d.json' % (n, i)))['d2h_pinned_gbs'] for i in range(n)]
print(('' if n == 1 else ',') + json.dumps({'gpus_copying': n, 'per_gpu_gbs': per, 'aggregate_gbs': sum(per),
      'what': 'mean over 12 x 1 GiB cudaMemcpyAsync D2H into pinned host memory, all GPUs at once'}))
PY
  rm -f gpurun_out/d2h_${n}_*.json gpurun_out/d2h_${n}_*.err
done
echo "]" >> gpurun_out/d2h_ceiling.json
cat gpurun_out/d2h_ceiling.json
if [ "$NG" -gt 1 ] && [[ " $* " != *" nobench "* ]]; then
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $NG --steps 3 --warmup 3 > gpurun_out/bench_n$NG.json 2> gpurun_out/bench_n$NG.err
  echo "bench n=$NG exit $?"; tail -c 2500 gpurun_out/bench_n$NG.json
fi
