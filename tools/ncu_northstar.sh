#!/bin/bash
# The three ncu counters BASELINE.json's north_star names, for every DAU kernel of one step of a layer:
# achieved HBM bytes/s, FP32 FMA pipe utilisation, shared-memory bank-conflict rate.
# usage: tools/ncu_northstar.sh <tag> [quick_time.py shape arguments]      (default shape: C2)
# Run on the GPU box; the same command runs first WITHOUT ncu.  Summarise with tools/ncu_northstar.py.
TAG=${1:-c2}; shift
mkdir -p gpurun_out
python tools/quick_time.py --iters 2 "$@" > gpurun_out/quick_time_$TAG.log 2>&1 || exit 1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed
M=$M,sm__inst_executed_pipe_fma.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed
M=$M,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_elapsed
M=$M,sm__inst_executed_pipe_lsu.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_elapsed
M=$M,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
M=$M,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.max,sm__throughput.avg.pct_of_peak_sustained_elapsed
# the second iteration only: skip the kernels of the first one (their count depends on the shape: tail kernels)
K='dau_gather|dau_grad|blur_pair|error_relayout|grad_finalize|unit_prep|gather_tail'
N1=$(ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" --csv python tools/quick_time.py --iters 1 "$@" 2>/dev/null | grep -c '^"[0-9]')
ncu --metrics $M --clock-control none -k regex:"$K" -s $N1 -c $N1 --csv \
    --log-file gpurun_out/ncu_northstar_$TAG.csv python tools/quick_time.py --iters 2 "$@" > gpurun_out/ncu_northstar_$TAG.log 2>&1
python tools/ncu_northstar.py gpurun_out/ncu_northstar_$TAG.csv > gpurun_out/ncu_northstar_$TAG.txt 2>&1
tail -14 gpurun_out/ncu_northstar_$TAG.txt
