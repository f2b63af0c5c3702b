#!/bin/bash
# multi-GPU evidence run (gpurun --gpus 8): GPU tests on the 8-GPU box, TP peer test at N = 2 / 8, bench at N = 1 / 2 / 8
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 10 1200 python -m pytest tests -m gpu -x -q > $O/m_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/m_pytest.log
timeout -k 10 600 $TR --nproc-per-node 2 --master-port 29511 tests/gpu_tp_peer.py > $O/m_tp_peer_n2.log 2>&1; echo "tp n2 rc=$?"
timeout -k 10 600 $TR --nproc-per-node 8 --master-port 29512 tests/gpu_tp_peer.py > $O/m_tp_peer_n8.log 2>&1; echo "tp n8 rc=$?"
timeout -k 10 600 python bench.py > $O/m_bench_n1.json 2> $O/m_bench_n1.err; echo "bench n1 rc=$?"
for n in 2 4 8; do
  timeout -k 10 900 $TR --nproc-per-node $n --master-port $((29520 + n)) bench.py --gpus $n > $O/m_bench_n$n.json 2> $O/m_bench_n$n.err; echo "bench n$n rc=$?"
done
python tools/benchsum.py $O/m_bench_n*.json
