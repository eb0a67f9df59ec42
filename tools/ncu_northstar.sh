#!/bin/bash
# The three ncu counters BASELINE.json's north_star names, for every DAU kernel of one C2 step:
# achieved HBM bytes/s, FP32 FMA pipe utilisation, shared-memory bank-conflict rate.
# Run on the GPU box AFTER the same command has exited 0 without ncu; summarise with tools/ncu_northstar.py.
mkdir -p gpurun_out
python tools/quick_time.py --iters 2 > gpurun_out/quick_time_plain.log 2>&1 || exit 1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed
M=$M,sm__inst_executed_pipe_fma.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed
M=$M,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_elapsed
M=$M,sm__inst_executed_pipe_lsu.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_elapsed
M=$M,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
M=$M,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.max,sm__throughput.avg.pct_of_peak_sustained_elapsed
ncu --metrics $M --clock-control none -k regex:'dau_gather|dau_grad|blur_pair|error_relayout|grad_finalize|unit_prep|gather_tail' -s 9 -c 9 --csv \
    --log-file gpurun_out/ncu_northstar.csv python tools/quick_time.py --iters 2 > gpurun_out/ncu_northstar.log 2>&1
tail -3 gpurun_out/ncu_northstar.log
