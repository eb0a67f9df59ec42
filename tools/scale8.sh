#!/bin/bash
# 8-GPU (or $1-GPU) scaling runs: C2 layer bench (weak) and the AlexNet-DAU step (strong, global batch 256)
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
$TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c2_n$N.json 2> gpurun_out/bench_c2_n$N.err
$TR tools/alexnet_dau_train_step.py --variant default --global-batch 256 > gpurun_out/alexnet_default_n$N.json 2> gpurun_out/alexnet_default_n$N.err
$TR tools/alexnet_dau_train_step.py --variant large --global-batch 2048 > gpurun_out/alexnet_large_gb2048_n$N.json 2> gpurun_out/alexnet_large_n$N.err
python tools/alexnet_dau_train_step.py --variant default --global-batch 256 > gpurun_out/alexnet_default_n1.json 2>/dev/null
tail -n 2 gpurun_out/*_n$N.err | tail -20
cut -c1-330 gpurun_out/bench_c2_n$N.json gpurun_out/alexnet_default_n$N.json gpurun_out/alexnet_large_gb2048_n$N.json gpurun_out/alexnet_default_n1.json
