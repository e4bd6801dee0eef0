#!/bin/bash
# scaling runs on one multi-GPU box: N = 1,2,4,8 (as available) for the 4k / 1080p / 16k workloads.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "gpus: $NG"
PORT=29511
for wl in ${WLS:-4k 1080p 16k}; do
  for n in 1 2 4 8; do
    [ $n -gt $NG ] && continue
    extra=""
    [ $wl = 1080p ] && extra="--frames 128"
    out=gpurun_out/scale_${wl}_n${n}.log
    if [ $n = 1 ]; then
      timeout 600 python bench.py --gpus 1 --workload $wl --steps ${STEPS:-5} --warmup 3 --no-cpu $extra > $out 2> ${out%.log}.err
    else
      PORT=$((PORT+1))
      NCCL_DEBUG=${NCCL_DEBUG_LEVEL:-WARN} timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --gpus $n --workload $wl --steps ${STEPS:-5} --warmup 3 --no-cpu $extra > $out 2> ${out%.log}.err
    fi
    echo "$wl n=$n exit $?"
    grep -o '"value": [0-9.]*, "unit": "MP/s", "n_gpus": [0-9]*' $out | head -1
    grep -o '"e2e": {"value": [0-9.]*' $out | head -1
    grep -o '"phase_ms_max_over_ranks": {[^}]*}' $out | head -1
  done
done
