#!/bin/bash
cd "$(dirname "$0")/.."
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 10 300 $TR --nproc-per-node 2 --master-port 29631 tools/soak_tp.py --launches 100000 > gpurun_out/soak_tp_n2.log 2>&1; echo "n2 rc=$?"
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29632 tools/soak_tp.py --launches 100000 > gpurun_out/soak_tp_n8.log 2>&1; echo "n8 rc=$?"
grep -h "^{" gpurun_out/soak_tp_n2.log gpurun_out/soak_tp_n8.log
