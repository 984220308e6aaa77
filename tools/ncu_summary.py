"""Print the handful of ncu raw-page metrics this project tracks: python tools/ncu_summary.py report.ncu-rep [row]"""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__block_size', 'launch__grid_size',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__cycles_active.avg', 'sm__cycles_elapsed.max',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum']
for r in rows[2:]:
    print('-' * 100)
    for k in keys:
        if k in h:
            i = h.index(k)
            print('%-70s %18s %s' % (k, r[i][:60], units[i]))
    for i, name in enumerate(h):
        if name.startswith('smsp__average_warps_issue_stalled') and name.endswith('per_issue_active.ratio'):
            v = float(r[i] or 0)
            if v >= 0.05:
                print('  stall %-60s %6.3f' % (name[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')], v))
