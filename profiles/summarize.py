#!/usr/bin/env python
"""Turn an .ncu-rep (brought back in gpurun_out/) into the text summary kept here.

    python profiles/summarize.py gpurun_out/prof_r1b.ncu-rep profiles/r1b_k_score.txt
"""
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__block_size', 'launch__grid_size',
        'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']


def main(rep, out_path, top=25):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    lines = ['ncu --set full --clock-control none --import-source on; source: %s' % rep, '']
    name = dict(zip(hdr, vals)).get('Kernel Name', '?')
    lines.append('kernel: %s' % name)
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS:
            lines.append('%-62s %14s %s' % (h, v, u))
    for h, u, v in zip(hdr, units, vals):
        if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio'):
            try:
                if float(v) >= 0.3:
                    lines.append('%-62s %14s %s' % (h.replace('smsp__average_warps_issue_stalled_', 'stall/'), v, u))
            except ValueError:
                pass
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'],
                         capture_output=True, text=True).stdout
    cur, head, acc = None, None, []
    for r in csv.reader(src.splitlines()):
        if r and r[0] == 'File Path':
            cur = r[1].split('/')[-1]
        elif r and r[0] == 'Line No':
            head = r
        elif r and head and r[0] not in ('', 'Function Name'):
            d = dict(zip(head, r))
            try:
                acc.append((cur, int(r[0]), r[1][:96], int(r[4] or 0), int(d.get('Instructions Executed') or 0)))
            except ValueError:
                pass
    ts = sum(a[3] for a in acc) or 1
    ti = sum(a[4] for a in acc) or 1
    lines += ['', 'top source lines by stall samples (%% of samples, %% of warp instructions)']
    for a in sorted(acc, key=lambda x: -x[3])[:top]:
        lines.append('%-16s %4d %6.2f%% %6.2f%%  %s' % (a[0][:16], a[1], 100 * a[3] / ts, 100 * a[4] / ti, a[2]))
    lines += ['', 'top source lines by executed warp instructions']
    for a in sorted(acc, key=lambda x: -x[4])[:top]:
        lines.append('%-16s %4d %6.2f%% %6.2f%%  %s' % (a[0][:16], a[1], 100 * a[3] / ts, 100 * a[4] / ti, a[2]))
    open(out_path, 'w').write('\n'.join(lines) + '\n')
    print('\n'.join(lines[:40]))


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2])
