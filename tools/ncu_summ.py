import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
want = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__warps_eligible.avg.per_cycle_active', 'smsp__warps_active.avg.per_cycle_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'sm__cycles_elapsed.avg', 'launch__grid_size', 'launch__block_size']
for r in rows[2:]:
    print('---', r[idx['Kernel Name']][:60])
    for w in want:
        if w in idx: print('  ', w, r[idx[w]], rows[1][idx[w]])
    st = []
    for h in hdr:
        if 'issue_stalled' in h and 'per_warp_active' in h and 'not_issued' not in h:
            try:
                v = float(r[idx[h]])
                if v > 1.0: st.append((v, h.split('issue_stalled_')[1].split('_per_warp')[0]))
            except ValueError: pass
    print('   stalls(%):', sorted(st, reverse=True))
