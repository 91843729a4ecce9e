"""Print the key metrics of every kernel in an `ncu --page raw --csv` dump: python tools/ncu_metrics.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum',
        'smsp__inst_executed.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.max']
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(f"{w} [{units[i]}]", [r[i][:42] for r in rows[2:]])
for w in hdr:
    if 'stalled' in w and 'per_issue_active' in w and 'not_issued' not in w:
        i = hdr.index(w)
        vals = [r[i][:6] for r in rows[2:]]
        if any(float(v or 0) > 0.2 for v in vals):
            print('stall', w.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), vals)
