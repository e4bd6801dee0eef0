#!/bin/bash
# Multi-GPU measurements on ONE box (gpurun --gpus 8), parameterised; every line lands in gpurun_out/.
#   PCIE="1 8"        ranks for tools/pcie_bw.py (raw pinned-copy ceiling of the e2e leg); PCIE_AFF=1 adds an --affinity run at the largest N
#   RUNS="wl:N:extra;..."   bench.py runs, e.g. "4k:8:;1080p:8:--frames 128;16k:8:;16k:8:--bins 1024"
mkdir -p gpurun_out
TAG=${TAG:-r02}
NG=$(nvidia-smi -L | wc -l)
echo "gpus: $NG"
( lscpu | egrep 'Model name|^CPU\(s\)|NUMA|Socket|Thread'; free -g | head -2; nvidia-smi topo -m ) > gpurun_out/host_${TAG}.txt 2>&1
PORT=29611
launch() {  # launch N script args...
  local n=$1; shift
  if [ "$n" = 1 ]; then python "$@"; else
    PORT=$((PORT+1))
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@"
  fi
}
last=1
for n in ${PCIE:-}; do
  [ $n -gt $NG ] && continue
  last=$n
  timeout 300 bash -c "$(declare -f launch); PORT=$PORT; launch $n tools/pcie_bw.py" > gpurun_out/pcie_${TAG}_n$n.json 2> gpurun_out/pcie_${TAG}_n$n.err
  echo "pcie n=$n exit $?"; cat gpurun_out/pcie_${TAG}_n$n.json | cut -c1-600
  PORT=$((PORT+3))
done
if [ "${PCIE_AFF:-0}" = 1 ] && [ $last -gt 1 ]; then
  timeout 300 bash -c "$(declare -f launch); PORT=$PORT; launch $last tools/pcie_bw.py --affinity" > gpurun_out/pcie_${TAG}_n${last}_aff.json 2> gpurun_out/pcie_${TAG}_n${last}_aff.err
  echo "pcie affinity n=$last exit $?"; cat gpurun_out/pcie_${TAG}_n${last}_aff.json | cut -c1-600
  PORT=$((PORT+3))
fi
i=0
IFS=';' read -ra RL <<< "${RUNS:-}"
for spec in "${RL[@]}"; do
  [ -z "$spec" ] && continue
  i=$((i+1))
  wl=${spec%%:*}; rest=${spec#*:}; n=${rest%%:*}; extra=${rest#*:}
  [ $n -gt $NG ] && continue
  out=gpurun_out/scale_${TAG}_${i}_${wl}_n${n}.log
  PORT=$((PORT+3))
  NCCL_DEBUG=${NCCL_DEBUG_LEVEL:-WARN} timeout 600 bash -c "$(declare -f launch); PORT=$PORT; launch $n bench.py --gpus $n --workload $wl --steps ${STEPS:-5} --warmup 3 --no-cpu $extra" > $out 2> ${out%.log}.err
  echo "[$wl n=$n $extra] exit $?"
  grep -o '"value": [0-9.]*, "unit": "MP/s", "n_gpus": [0-9]*' $out | head -1
  grep -o '"e2e": {"value": [0-9.]*' $out | head -1
  grep -o '"phase_ms_max_over_ranks": {[^}]*}' $out | head -1; grep -o '"host_enqueue": {[^}]*}' $out | head -1
  tail -2 ${out%.log}.err | cut -c1-300
done
