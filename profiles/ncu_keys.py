"""Print the key raw metrics of every kernel in an ncu report (run: ncu -i X --page raw --csv | python ncu_keys.py)."""
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[0]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'smsp__inst_executed.sum',
        'sass__thread_inst_executed_true_per_opcode', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed_op_shared_atom.sum',
        'smsp__inst_executed_op_branch.sum', 'launch__waves_per_multiprocessor', 'sm__cycles_elapsed.max']
# every pipe-utilisation counter the report has (DESIGN.md section 4 argues from these)
want += [hn for hn in hdr if hn not in want and (hn.startswith('sm__pipe_') and hn.endswith('pct_of_peak_sustained_active')
         or hn.startswith('sm__inst_executed_pipe_') and hn.endswith('.sum'))]
for r in rows[2:]:
    print('--', r[hdr.index('Kernel Name')][:60])
    for w in want:
        if w in hdr:
            print('   %-70s %s %s' % (w, r[hdr.index(w)], rows[1][hdr.index(w)]))
    for i, hn in enumerate(hdr):
        if 'issue_stalled' in hn and hn.endswith('per_issue_active.ratio'):
            try:
                v = float(r[i])
            except ValueError:
                continue
            if v > 0.3:
                print('   stall %-50s %.2f' % (hn.split('issue_stalled_')[1].split('_per_')[0], v))
