"""Summarise an `ncu --page raw --csv` dump: the metrics the DAU kernels are judged on."""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__inst_executed_pipe_fma.sum.pct', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy', 'sm__pipe_fmalite', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'lts__t_bytes.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max', 'sm__throughput.avg.pct',
        'smsp__average_warps_issue_stalled', 'lts__t_sector_hit_rate', 'sm__inst_executed_pipe_lsu',
        'smsp__inst_executed_op_shared_ld', 'sm__inst_executed_pipe_alu', 'sm__inst_executed_pipe_fp']
skip = ['per_second', 'peak_sustained', '.min', '.max.', 'pcsamp']
idx = [i for i, h in enumerate(hdr) if any(w in h for w in want) and not any(s in h for s in skip)]
for r in rows[2:]:
    print('-----')
    for i in idx:
        v = r[i]
        if 'stalled' in hdr[i]:
            try:
                if float(v) < 0.05: continue
            except ValueError:
                pass
        print(hdr[i], '=', v, units[i])
