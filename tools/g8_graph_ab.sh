cd /root/repo
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
for i in 1 2; do
timeout -k 10 300 $TR --master-port 2961$i bench.py --gpus 8 --no-extra > gpurun_out/g8_eager_$i.json 2> gpurun_out/g8_eager_$i.err; echo "eager rc=$?"
timeout -k 10 300 $TR --master-port 2962$i bench.py --gpus 8 --no-extra --graph > gpurun_out/g8_graph_$i.json 2> gpurun_out/g8_graph_$i.err; echo "graph rc=$?"
done
for f in gpurun_out/g8_*.json; do grep "^{" $f | tail -1 > $f.line; done
python tools/benchsum.py gpurun_out/g8_*.json.line
