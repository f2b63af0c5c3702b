#!/bin/bash
# round-2 evidence run (1 GPU): default bench line, reference arm, ncu launch list + full captures
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
( time python bench.py > $O/p_bench_default.json 2> $O/p_bench_default.err ) 2> $O/p_bench_default.time; echo "bench rc=$?"
( time python bench.py --impl reference > $O/p_bench_reference.json 2> $O/p_bench_reference.err ) 2> $O/p_bench_reference.time; echo "ref rc=$?"
Q="--steps 3 --warmup 3 --no-extra --no-cpu-baseline"
python bench.py $Q > $O/p_plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spa_|FillFunctor|reduce_kernel" -c 60 --csv --log-file $O/p_launches_c2.csv python bench.py $Q > $O/p_ncu_launches.log 2>&1
echo "launch list rc=$?"
for wl in c2 c3 c4; do
  python bench.py $Q --workload $wl > $O/p_plain_$wl.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:spa_chain -s 4 -c 1 -f -o $O/p_full_$wl python bench.py $Q --workload $wl > $O/p_ncu_full_$wl.log 2>&1
  echo "ncu full $wl rc=$?"
done
ls -la $O | grep p_
for wl in c2 c3 c4; do timeout -k 10 300 python tests/gpu_phase_trace.py $wl > $O/p_trace_$wl.txt 2>&1; done
timeout -k 10 900 python tools/sweep.py > $O/p_sweep.md 2> $O/p_sweep.err; echo "sweep rc=$?"
timeout -k 10 300 python tools/bench_assign.py > $O/p_assign.json 2>&1; echo "assign rc=$?"
