"""Brief per-launch table of an .ncu-rep: python tools/ncu_brief.py file.ncu-rep"""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'lts__t_bytes.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'l1tex__t_bytes.sum', 'sm__sass_inst_executed_op_shared.sum']
ki = h.index('Kernel Name')
for r in rows[2:]:
    print('==', r[ki][:110])
    for w in want:
        if w in h:
            i = h.index(w)
            print('   %-70s %s %s' % (w, r[i], u[i]))
    st = sorted(((float(r[i] or 0), h[i]) for i in range(len(h)) if 'pcsamp_warps_issue_stalled' in h[i] and 'not_issued' not in h[i]), reverse=True)
    print('   stalls:', ', '.join('%s %d' % (n.split('issue_stalled_')[1], v) for v, n in st[:7]))
