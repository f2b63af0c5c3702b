#!/bin/bash
# round-2 GPU batch 1: smoke, parity suite, bench A/B of the new schedule, phase traces
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > $O/b1_gpu.txt 2>&1
timeout -k 10 180 python __graft_entry__.py --smoke > $O/b1_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/b1_status.txt
timeout -k 10 900 python -m pytest tests -m gpu -x -q > $O/b1_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/b1_status.txt
tail -5 $O/b1_pytest.log
for wl in c2 c3 c4; do
  timeout -k 10 300 python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline > $O/b1_bench_$wl.json 2> $O/b1_bench_$wl.err; echo "bench $wl rc=$?" | tee -a $O/b1_status.txt
done
SPA_COOP=0 timeout -k 10 300 python bench.py --workload c2 --steps 30 --warmup 5 --no-cpu-baseline > $O/b1_bench_c2_nocoop.json 2>&1
for dr in ; do
  SPA_DYN_RUNS=$dr timeout -k 10 300 python bench.py --workload c2 --steps 30 --warmup 5 --no-cpu-baseline > $O/b1_bench_c2_dr$dr.json 2>&1
  SPA_DYN_RUNS=$dr timeout -k 10 300 python bench.py --workload c3 --steps 30 --warmup 5 --no-cpu-baseline > $O/b1_bench_c3_dr$dr.json 2>&1
done
for sp in 35 65; do
  SPA_STATIC_PCT=$sp timeout -k 10 300 python bench.py --workload c2 --steps 30 --warmup 5 --no-cpu-baseline > $O/b1_bench_c2_sp$sp.json 2>&1
done
for wl in c2 c3 c4; do
  timeout -k 10 300 python tests/gpu_phase_trace.py $wl > $O/b1_trace_$wl.txt 2>&1
done
python tools/benchsum.py $O/b1_bench_*.json 2>/dev/null | tee $O/b1_summary.txt
