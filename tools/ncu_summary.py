#!/usr/bin/env python
"""Prints the metrics we track from an ncu report (`ncu -i X.ncu-rep --page raw --csv` piped in or given as a path)."""
import csv, subprocess, sys
WANT = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.avg.per_second',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__warps_eligible.avg.per_cycle_active',
        'sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_tensor.sum',
        'l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum', 'l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'sm__cycles_active.avg']
STALL = 'smsp__average_warps_issue_stalled_'
def main(path):
    txt = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('--- ' + r[hdr.index('Kernel Name')][:90])
        for w in WANT:
            cand = [h for h in hdr if h == w or h.endswith('.' + w)]
            if w not in hdr and cand:
                print(f'  {w} = {r[hdr.index(cand[0])]} {units[hdr.index(cand[0])]}')
            if w in hdr:
                print(f'  {w} = {r[hdr.index(w)]} {units[hdr.index(w)]}')
        st = [(float(r[i] or 0), hdr[i][len(STALL):].replace('_per_issue_active.ratio', '')) for i in range(len(hdr))
              if hdr[i].startswith(STALL) and hdr[i].endswith('_per_issue_active.ratio')]
        print('  stalls/issue: ' + ', '.join(f'{n}={v:.2f}' for v, n in sorted(st, reverse=True)[:8]))
if __name__ == '__main__':
    main(sys.argv[1])
