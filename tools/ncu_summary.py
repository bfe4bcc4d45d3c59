#!/usr/bin/env python
"""Summarise an .ncu-rep: key metrics per kernel (raw page) and the instruction distribution over the SASS of one
kernel (source page).  Usage: ncu_summary.py report.ncu-rep [kernel-regex-for-source-page] [block]"""
import csv
import io
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'smsp__issue_active.avg.pct',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp16.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__inst_executed_op_shared_atom.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'sm__cycles_active.avg', 'smsp__cycles_active.avg']


def page(rep, name, extra=()):
    out = subprocess.run(['ncu', '-i', rep, '--page', name, '--csv', *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = page(rep, 'raw')
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('====', r[hdr.index('Kernel Name')][:90])
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f'  {w:75s} {r[i]} {units[i]}')
        for i, h in enumerate(hdr):
            if 'smsp__average_warps_issue_stalled' in h and '_per_issue_active' in h:
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                if v > 0.3:
                    print(f'    stall {h.split("issue_stalled_")[1].split("_per_issue")[0]:28s} {v:.2f}')
    if len(sys.argv) > 2:
        blk = int(sys.argv[3]) if len(sys.argv) > 3 else 40
        rows = page(rep, 'source', ['--kernel-name', 'regex:' + sys.argv[2]])
        hdr = rows[1]
        ia, isrc, ie, isamp, it = (hdr.index(n) for n in ('Address', 'Source', 'Instructions Executed', '# Samples', 'Avg. Threads Executed'))
        data = []
        for r in rows[2:]:
            try:
                data.append((int(r[ia], 16), r[isrc].strip(), int(r[ie]), int(r[isamp]), r[it]))
            except (ValueError, IndexError):
                break                      # next kernel's header
        base = data[0][0]
        tot = sum(d[2] for d in data)
        ts = sum(d[3] for d in data)
        print(f'---- source page: {len(data)} SASS instructions, {tot} executed, {ts} samples')
        for i in range(0, len(data), blk):
            b = data[i:i + blk]
            s = sum(d[2] for d in b)
            sm = sum(d[3] for d in b)
            print(f'{b[0][0] - base:6x} {100 * s / tot:5.1f}% inst {100 * sm / max(ts, 1):5.1f}% samples  thr {b[0][4]:>5s}  {b[0][1][:70]}')


if __name__ == '__main__':
    main()
