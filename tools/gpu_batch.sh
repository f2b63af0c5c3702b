#!/bin/bash
# generic GPU batch: tools/gpu_batch.sh TAG [variant ...]   variant = name:wl[,wl..]:ENV=V[,ENV=V..]
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O; T=$1; shift
timeout -k 10 180 python __graft_entry__.py --smoke > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee $O/${T}_status.txt
if [ -z "$SKIP_TESTS" ]; then
timeout -k 10 1200 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_status.txt
tail -5 $O/${T}_pytest.log
fi
for v in "$@"; do
  name=${v%%:*}; rest=${v#*:}; wls=${rest%%:*}; envs=${rest#*:}
  [ "$envs" == "$rest" ] && envs="X=1"
  for wl in ${wls//,/ }; do
    env ${envs//,/ } timeout -k 10 300 python bench.py --workload $wl --steps ${STEPS:-30} --warmup 5 --no-cpu-baseline > $O/${T}_bench_${wl}_$name.json 2> $O/${T}_bench_${wl}_$name.err
  done
done
for wl in ${TRACE:-c2 c3 c4}; do
  timeout -k 10 300 python tests/gpu_phase_trace.py $wl > $O/${T}_trace_$wl.txt 2>&1
done
python tools/benchsum.py $O/${T}_bench_*.json 2>/dev/null | tee $O/${T}_summary.txt
