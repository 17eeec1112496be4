"""Summarise an .ncu-rep (read on the CPU box): key roofline metrics per kernel + stall mix.  usage: ncu_summary.py rep"""
import csv, subprocess, sys, io, re
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'sm__cycles_elapsed.max', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__warps_eligible.avg.per_cycle_active']
for r in rows[2:]:
    print('====', r[hdr.index('Kernel Name')][:90])
    for k in keys:
        if k in hdr:
            i = hdr.index(k); print('  %-85s %s %s' % (k, r[i], units[i]))
    st = [(h, float(r[i])) for i, h in enumerate(hdr) if re.match(r'smsp__average_warps_issue_stalled_.*_per_issue_active.ratio', h) and r[i]]
    st.sort(key=lambda x: -x[1])
    print('  stalls/issue:', ', '.join('%s %.2f' % (h.split('stalled_')[1].split('_per_issue')[0], v) for h, v in st[:8]))
