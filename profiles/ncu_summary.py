#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here with `ncu -i`, no GPU needed): per-launch duration, DRAM bytes, pipe
utilisation, occupancy and the top stall reasons.  Usage: python profiles/ncu_summary.py gpurun_out/x.ncu-rep"""
import csv
import io
import subprocess
import sys

KEYS = [
    ('gpu__time_duration.sum', 'duration'),
    ('dram__bytes_read.sum', 'dram read'),
    ('dram__bytes_write.sum', 'dram write'),
    ('dram__throughput.avg.pct_of_peak_sustained_elapsed', 'dram % of peak'),
    ('lts__t_bytes.sum', 'L2 bytes'),
    ('launch__registers_per_thread', 'regs/thread'),
    ('launch__occupancy_limit_registers', 'occ limit regs (blocks)'),
    ('launch__waves_per_multiprocessor', 'waves/SM'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'achieved occupancy %'),
    ('smsp__inst_executed.sum', 'warp instructions'),
    ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue slots busy %'),
    ('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'fp64 pipe %'),
    ('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'alu pipe %'),
    ('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'fma pipe %'),
    ('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'lsu pipe %'),
    ('sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'xu pipe %'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'SM throughput %'),
    ('l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct', 'L1 load hit %'),
    ('smsp__thread_inst_executed_per_inst_executed.ratio', 'active threads/instr'),
]


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print('kernel:', data[0][col['Kernel Name']][:100])
    print('grid/block:', data[0][col['Grid Size']], data[0][col['Block Size']])
    for k, name in KEYS:
        if k in col:
            print('%-28s %-8s %s' % (name, units[col[k]], '  '.join(r[col[k]] for r in data)))
    stalls = [(h, i) for h, i in col.items() if 'average_warp_latency_issue_stalled' in h and h.endswith('.ratio')
              or ('warp_issue_stalled' in h and h.endswith('_per_warp_active.pct'))]
    vals = []
    for h, i in stalls:
        try:
            vals.append((float(data[-1][i]), h))
        except ValueError:
            pass
    print('top stall reasons (last launch):')
    for v, h in sorted(vals, reverse=True)[:8]:
        print('   %8.2f  %s' % (v, h.replace('smsp__average_warp_latency_issue_stalled_', '').replace('smsp__warp_issue_stalled_', '')))


if __name__ == '__main__':
    main(sys.argv[1])
