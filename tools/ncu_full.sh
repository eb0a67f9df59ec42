#!/bin/bash
# One `ncu --set full` capture of the hot kernels of a layer step (after the same command ran without ncu):
# traffic (dram bytes), stall reasons, source page.  usage: tools/ncu_full.sh <tag> [quick_time.py shape args]
TAG=${1:-c2}; shift
python tools/quick_time.py --iters 1 "$@" > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'dau_(grad|gather)_tma_kernel' -c 3 -f -o gpurun_out/full_$TAG \
    python tools/quick_time.py --iters 1 "$@" > gpurun_out/ncu_full_$TAG.log 2>&1
ncu -i gpurun_out/full_$TAG.ncu-rep --page raw --csv > gpurun_out/full_${TAG}_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/full_${TAG}_raw.csv > gpurun_out/full_${TAG}_summary.txt 2>&1
grep -E "Kernel Name|dram__bytes_(read|write).sum =|gpu__time_duration.sum" gpurun_out/full_${TAG}_summary.txt | head -12
