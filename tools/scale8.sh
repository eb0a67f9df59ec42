#!/bin/bash
# N-GPU session (default 8): host copy ceiling, C2 layer bench (weak) with the AlexNet-DAU block (strong, global
# batch 256, eager and graphed), and the AlexNet-DAU step alone.  Outputs under gpurun_out/ with the given tag.
N=${1:-8}
TAG=${2:-r2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for opt in "" "--bind 1"; do
  timeout 120 $TR tools/pcie_probe.py $opt 2>/dev/null | grep '^{' >> gpurun_out/${TAG}_pcie_probe_n$N.jsonl
done
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 2> gpurun_out/${TAG}_bench_c2_n$N.err | grep '^{' > gpurun_out/${TAG}_bench_c2_n$N.json
timeout 200 $TR tools/alexnet_dau_train_step.py --variant default --global-batch 256 --graph 0 2> gpurun_out/${TAG}_alexnet_n$N.err | grep '^{' > gpurun_out/${TAG}_alexnet_default_eager_n$N.json
timeout 200 $TR tools/alexnet_dau_train_step.py --variant default --global-batch 256 --graph 1 2>> gpurun_out/${TAG}_alexnet_n$N.err | grep '^{' > gpurun_out/${TAG}_alexnet_default_graph_n$N.json
python tools/alexnet_dau_train_step.py --variant default --global-batch 256 --graph 1 2>/dev/null | grep '^{' > gpurun_out/${TAG}_alexnet_default_graph_n1.json
tail -n 3 gpurun_out/${TAG}_*_n$N.err | tail -12
cut -c1-420 gpurun_out/${TAG}_pcie_probe_n$N.jsonl gpurun_out/${TAG}_alexnet_default_*_n$N.json gpurun_out/${TAG}_alexnet_default_graph_n1.json
