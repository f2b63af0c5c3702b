#!/bin/bash
# round-2 GPU batch 2: parity suite + schedule A/B (late claims, guided chunks)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O; T=b2
timeout -k 10 180 python __graft_entry__.py --smoke > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee $O/${T}_status.txt
timeout -k 10 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_status.txt
tail -5 $O/${T}_pytest.log
run() { # name workload env...
  local name=$1 wl=$2; shift 2
  env "$@" timeout -k 10 300 python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline > $O/${T}_bench_${wl}_$name.json 2> $O/${T}_bench_${wl}_$name.err
}
for wl in c2 c3 c4; do run base $wl X=1; done
for wl in c2 c3; do
  run lead2 $wl SPA_CLAIM_LEAD=2
  run lead8 $wl SPA_CLAIM_LEAD=8
  run f13 $wl SPA_CHUNK_FRAC=1/3
  run sp35 $wl SPA_STATIC_PCT=35
  run sp65 $wl SPA_STATIC_PCT=65
  run sp80 $wl SPA_STATIC_PCT=80
done
run sp65 c4 SPA_STATIC_PCT=65
run sp80 c4 SPA_STATIC_PCT=80
for wl in c2 c3 c4; do
  timeout -k 10 300 python tests/gpu_phase_trace.py $wl > $O/${T}_trace_$wl.txt 2>&1
done
python tools/benchsum.py $O/${T}_bench_*.json 2>/dev/null | tee $O/${T}_summary.txt
